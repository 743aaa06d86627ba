// Host-side (setup-time) model of loudspeaker layouts and the construction of the
// device tables.  Runs once per calculator; nothing here is on the per-object path.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "tables.h"

namespace earb {

  // error categories mirror libear's five exception classes
  // (ref include/ear/exceptions.hpp:8-43); the C ABI returns them as status codes.
  enum Status {
    OK = 0,
    NOT_IMPLEMENTED = 1,
    INTERNAL_ERROR = 2,
    ADM_ERROR = 3,
    UNKNOWN_LAYOUT = 4,
    INVALID_ARGUMENT = 5,
    CUDA_ERROR = 6,
  };

  struct Error : std::runtime_error {
    Status status;
    Error(Status s, const std::string &what) : std::runtime_error(what), status(s) {}
  };

  struct Polar {
    double az, el, dist;
  };

  struct Vec3 {
    double x, y, z;
  };

  struct HostChannel {
    std::string name;
    Polar pos;      // real position
    Polar nominal;  // nominal position
    double az_lo, az_hi, el_lo, el_hi;
    bool lfe;
  };

  struct HostLayout {
    std::string name;
    std::vector<HostChannel> channels;

    HostLayout withoutLfe() const;
    int indexOf(const std::string &channel_name) const;  // -1 if absent
  };

  // the tabulated ITU-R BS.2051 layouts (ref src/bs2051.cpp:13-22)
  const std::vector<std::string> &layoutNames();
  HostLayout getLayout(const std::string &name);  // throws UNKNOWN_LAYOUT

  Vec3 cart(double az, double el, double dist);  // ref src/common/geom.cpp:82-87

  // extra loudspeakers + A x S downmix (ref point_source_panner.cpp:258-340)
  struct ExtraSpeakers {
    std::vector<HostChannel> channels;
    std::vector<int> parent;  // parent[a] for a in [0, S + n_extra): real channel index
  };
  ExtraSpeakers extraPosVerticalNominal(const HostLayout &layout_no_lfe);

  // order of the vertices of a convex, roughly planar polygon (ref geom.cpp:38-69)
  std::vector<int> ngonVertexOrder(const std::vector<Vec3> &verts);

  // facets (sorted vertex index lists) of the convex hull of unit vectors, for layouts without a
  // tabulated hull (ref src/common/convex_hull.cpp:7-100)
  std::vector<std::vector<int>> convexHull(const std::vector<Vec3> &positions, double tolerance = 1e-5);

  // build the point-source panner tables for a layout (LFE channels are dropped
  // first, as every libear calculator does: ref gain_calculator_objects.cpp:28)
  void buildPspTables(const HostLayout &layout, PspTables &out);

  // the even panning grid of the extent panner (ref polar_extent.cpp:15-39)
  std::vector<Vec3> extentGridPositions(int n_rows = 37);

  // spherical t-design used by the HOA AllRAD design (ref src/hoa/hoa.cpp:4-15)
  const std::vector<Vec3> &hoaDesignPoints();

}  // namespace earb

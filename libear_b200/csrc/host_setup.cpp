// Host-side setup: layout tables -> augmented layout -> region tables.
//
// Written from scratch against the behaviour of libear's
// src/common/point_source_panner.cpp / geom.cpp / layout.cpp (cited per function);
// the output is the flat PspTables consumed by the CUDA kernels instead of libear's
// tree of RegionHandler objects.
#include "host_setup.hpp"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <set>

namespace earb {

  namespace {
    struct ChannelSpec {
      const char *name;
      double az, el, az_lo, az_hi, el_lo, el_hi;
      int lfe;
    };
    struct LayoutSpec {
      const char *name;
      const ChannelSpec *channels;
      int n_channels;
      const int *facets;  // vertex lists, -1 after each facet, -2 at the end
    };
#include "bs2051_tables.inc"

    const double kPi = 3.141592653589793238462643383279502884;

    inline double radians(double d) { return d * (kPi / 180.0); }

    Vec3 operator-(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
    Vec3 operator+(Vec3 a, Vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
    double dot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
    Vec3 cross(Vec3 a, Vec3 b) {
      return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
    }

    const LayoutSpec *findSpec(const std::string &name) {
      for (int i = 0; i < kNumLayouts; i++)
        if (name == kLayouts[i].name) return &kLayouts[i];
      return nullptr;
    }

    bool hasFacetTable(const std::string &name) {
      const LayoutSpec *spec = findSpec(name);
      return spec && spec->facets;
    }

    std::vector<std::vector<int>> facetsFor(const std::string &name) {
      const LayoutSpec *spec = findSpec(name);
      if (!spec || !spec->facets) throw Error(INTERNAL_ERROR, "no facet table for layout " + name);
      std::vector<std::vector<int>> out;
      std::vector<int> cur;
      for (const int *p = spec->facets; *p != -2; ++p) {
        if (*p == -1) {
          out.push_back(cur);
          cur.clear();
        } else {
          cur.push_back(*p);
        }
      }
      return out;
    }

    // Inverse of a 3x3 matrix (rows r0..r2), adjugate form evaluated in long double so
    // the rounded result is the correctly rounded inverse to within an ulp.  libear
    // uses Eigen's LU on a dynamic matrix (ref point_source_panner.cpp:38); both are
    // within a few ulp of the exact inverse, far inside the 1e-11 acceptance band.
    void inverse3(const Vec3 rows[3], double out[9]) {
      long double m[3][3] = {{rows[0].x, rows[0].y, rows[0].z},
                             {rows[1].x, rows[1].y, rows[1].z},
                             {rows[2].x, rows[2].y, rows[2].z}};
      long double c00 = m[1][1] * m[2][2] - m[1][2] * m[2][1];
      long double c01 = m[1][2] * m[2][0] - m[1][0] * m[2][2];
      long double c02 = m[1][0] * m[2][1] - m[1][1] * m[2][0];
      long double det = m[0][0] * c00 + m[0][1] * c01 + m[0][2] * c02;
      if (det == 0.0L) throw Error(INTERNAL_ERROR, "singular loudspeaker triplet");
      long double inv[3][3];
      inv[0][0] = c00;
      inv[1][0] = c01;
      inv[2][0] = c02;
      inv[0][1] = m[0][2] * m[2][1] - m[0][1] * m[2][2];
      inv[1][1] = m[0][0] * m[2][2] - m[0][2] * m[2][0];
      inv[2][1] = m[0][1] * m[2][0] - m[0][0] * m[2][1];
      inv[0][2] = m[0][1] * m[1][2] - m[0][2] * m[1][1];
      inv[1][2] = m[0][2] * m[1][0] - m[0][0] * m[1][2];
      inv[2][2] = m[0][0] * m[1][1] - m[0][1] * m[1][0];
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) out[3 * i + j] = (double)(inv[i][j] / det);
    }

    // rows of the quadratic's coefficient matrix for one bilinear axis of a quad
    // (ref point_source_panner.cpp:132-145): poly(pos) = [v0.pos, v1.pos, v2.pos]
    void polyBasis(const Vec3 &a, const Vec3 &b, const Vec3 &c, const Vec3 &d,
                   double out[9]) {
      Vec3 v0 = cross(b - a, c - d);
      Vec3 v1 = cross(a, c - d) + cross(b - a, d);
      Vec3 v2 = cross(a, d);
      const Vec3 v[3] = {v0, v1, v2};
      for (int i = 0; i < 3; i++) {
        out[3 * i] = v[i].x;
        out[3 * i + 1] = v[i].y;
        out[3 * i + 2] = v[i].z;
      }
    }

    void fillTriplet(PspTriplet &t, const Vec3 &p0, const Vec3 &p1, const Vec3 &p2,
                     int c0, int c1, int c2) {
      const Vec3 rows[3] = {p0, p1, p2};
      inverse3(rows, t.basis);
      t.ch[0] = c0;
      t.ch[1] = c1;
      t.ch[2] = c2;
      t.pad = 0;
    }

    // M+SC / M-SC azimuth checks (ref point_source_panner.cpp:560-578)
    void checkScreenSpeakers(const HostLayout &layout) {
      for (const auto &ch : layout.channels) {
        if (ch.name == "M+SC" || ch.name == "M-SC") {
          double abs_az = std::abs(ch.pos.az);
          if (!((5.0 <= abs_az && abs_az < 25.0) || (35.0 <= abs_az && abs_az < 60.0)))
            throw Error(INVALID_ARGUMENT,
                        "M+SC or M-SC has azimuth not in the allowed ranges of 5 to 25 "
                        "and 35 to 60 degrees");
          if (25.0 < abs_az)
            throw Error(NOT_IMPLEMENTED,
                        "not implemented: M+SC and M-SC with azimuths wider than 25 "
                        "degrees are not currently supported");
        }
      }
    }

    void buildFull(const HostLayout &layout, PspTables &T) {
      const int S = (int)layout.channels.size();
      if (S > EARB_MAX_SPK) throw Error(INVALID_ARGUMENT, "too many loudspeakers");

      // augmented layout (ref point_source_panner.cpp:425-471)
      ExtraSpeakers extra = extraPosVerticalNominal(layout);
      std::vector<Vec3> real;
      for (const auto &ch : layout.channels) real.push_back(cart(ch.pos.az, ch.pos.el, 1.0));
      for (const auto &ch : extra.channels) real.push_back(cart(ch.pos.az, ch.pos.el, 1.0));
      const int A = (int)real.size();
      if (A > EARB_MAX_AUG) throw Error(INVALID_ARGUMENT, "too many augmented loudspeakers");

      std::vector<int> virtual_verts;
      virtual_verts.push_back((int)real.size());
      real.push_back({0.0, 0.0, -1.0});
      if (layout.indexOf("T+000") < 0 && layout.indexOf("UH+180") < 0) {
        virtual_verts.push_back((int)real.size());
        real.push_back({0.0, 0.0, 1.0});
      }

      // hull facets: tabulated for the BS.2051 layout names, else the convex hull of the
      // *nominal* positions (ref point_source_panner.cpp:484-487)
      std::vector<std::vector<int>> facets;
      if (hasFacetTable(layout.name)) {
        facets = facetsFor(layout.name);
      } else {
        std::vector<Vec3> nominal;
        for (const auto &ch : layout.channels) nominal.push_back(cart(ch.nominal.az, ch.nominal.el, 1.0));
        for (const auto &ch : extra.channels) nominal.push_back(cart(ch.nominal.az, ch.nominal.el, 1.0));
        for (int vv : virtual_verts) nominal.push_back(real[vv]);
        facets = convexHull(nominal);
      }
      for (const auto &f : facets)
        for (int v : f)
          if (v < 0 || v >= (int)real.size())
            throw Error(INTERNAL_ERROR, "facet table does not match the layout");

      std::memset(&T, 0, sizeof(T));
      T.n_aug = A;
      T.n_real = S;
      T.stereo = 0;
      for (int a = 0; a < A; a++) T.parent[a] = extra.parent[a];

      int n_tri = 0, n_quad = 0, n_ngon = 0, n_regions = 0;
      auto is_virtual = [&](int v) {
        return std::find(virtual_verts.begin(), virtual_verts.end(), v) != virtual_verts.end();
      };

      // virtual n-gons first, in ascending virtual-vertex order (ref :495-518)
      for (int vv : virtual_verts) {
        std::set<int> adj;
        for (const auto &f : facets)
          if (std::find(f.begin(), f.end(), vv) != f.end()) adj.insert(f.begin(), f.end());
        adj.erase(vv);
        for (int v : adj)
          if (is_virtual(v)) throw Error(INTERNAL_ERROR, "invalid triangulation");
        std::vector<int> verts(adj.begin(), adj.end());
        const int n = (int)verts.size();
        if (n > EARB_MAX_NGON_VERTS || n_ngon >= EARB_MAX_NGON || n_tri + n > EARB_MAX_TRI)
          throw Error(INTERNAL_ERROR, "virtual n-gon too large");
        std::vector<Vec3> pos(n);
        for (int i = 0; i < n; i++) pos[i] = real[verts[i]];
        std::vector<int> order = ngonVertexOrder(pos);

        PspNgon &g = T.ngon[n_ngon];
        g.n = n;
        g.first_tri = n_tri;
        g.centre_downmix = 1.0 / std::sqrt((double)n);
        for (int i = 0; i < n; i++) g.ch[i] = verts[i];
        // inner triplets: (edge vertex i, edge vertex i+1, virtual centre) (ref :71-81)
        for (int i = 0; i < n; i++) {
          int j = (i + 1) % n;
          fillTriplet(T.tri[n_tri++], pos[order[i]], pos[order[j]], real[vv], order[i],
                      order[j], n);
        }
        T.regions[n_regions].kind = EARB_REGION_NGON;
        T.regions[n_regions].index = n_ngon++;
        n_regions++;
      }

      // remaining facets in table order -> triplets and quads (ref :522-554)
      for (const auto &f : facets) {
        bool touches_virtual = false;
        for (int v : f) touches_virtual |= is_virtual(v);
        if (touches_virtual) continue;
        if (n_regions >= EARB_MAX_REGIONS) throw Error(INTERNAL_ERROR, "too many regions");
        if (f.size() == 3) {
          if (n_tri >= EARB_MAX_TRI) throw Error(INTERNAL_ERROR, "too many triplets");
          T.regions[n_regions].kind = EARB_REGION_TRIPLET;
          T.regions[n_regions].index = n_tri;
          fillTriplet(T.tri[n_tri++], real[f[0]], real[f[1]], real[f[2]], f[0], f[1], f[2]);
        } else if (f.size() == 4) {
          if (n_quad >= EARB_MAX_QUAD) throw Error(INTERNAL_ERROR, "too many quads");
          PspQuad &q = T.quad[n_quad];
          std::vector<Vec3> pos(4);
          for (int i = 0; i < 4; i++) {
            pos[i] = real[f[i]];
            q.pos[3 * i] = pos[i].x;
            q.pos[3 * i + 1] = pos[i].y;
            q.pos[3 * i + 2] = pos[i].z;
            q.ch[i] = f[i];
          }
          std::vector<int> order = ngonVertexOrder(pos);
          for (int i = 0; i < 4; i++) q.order[i] = order[i];
          const Vec3 &a = pos[order[0]], &b = pos[order[1]], &c = pos[order[2]],
                     &d = pos[order[3]];
          polyBasis(a, b, c, d, q.polyX);  // ref :109
          polyBasis(b, c, d, a, q.polyY);  // vertices shifted by one (ref :107-110)
          T.regions[n_regions].kind = EARB_REGION_QUAD;
          T.regions[n_regions].index = n_quad++;
        } else {
          throw Error(INTERNAL_ERROR, "facets with more than 4 vertices are not supported");
        }
        n_regions++;
      }
      T.n_regions = n_regions;
    }
  }  // namespace

  Vec3 cart(double az, double el, double dist) {
    return {std::sin(radians(-az)) * std::cos(radians(el)) * dist,
            std::cos(radians(-az)) * std::cos(radians(el)) * dist,
            std::sin(radians(el)) * dist};
  }

  HostLayout HostLayout::withoutLfe() const {
    HostLayout out;
    out.name = name;
    for (const auto &c : channels)
      if (!c.lfe) out.channels.push_back(c);
    return out;
  }

  int HostLayout::indexOf(const std::string &channel_name) const {
    for (size_t i = 0; i < channels.size(); i++)
      if (channels[i].name == channel_name) return (int)i;
    return -1;
  }

  const std::vector<std::string> &layoutNames() {
    static const std::vector<std::string> names = [] {
      std::vector<std::string> n;
      for (int i = 0; i < kNumLayouts; i++) n.push_back(kLayouts[i].name);
      return n;
    }();
    return names;
  }

  HostLayout getLayout(const std::string &name) {
    const LayoutSpec *spec = findSpec(name);
    if (!spec) throw Error(UNKNOWN_LAYOUT, "unknown layout: " + name);
    HostLayout l;
    l.name = name;
    for (int i = 0; i < spec->n_channels; i++) {
      const ChannelSpec &c = spec->channels[i];
      HostChannel h;
      h.name = c.name;
      h.pos = {c.az, c.el, 1.0};
      h.nominal = {c.az, c.el, 1.0};
      h.az_lo = c.az_lo;
      h.az_hi = c.az_hi;
      h.el_lo = c.el_lo;
      h.el_hi = c.el_hi;
      h.lfe = c.lfe != 0;
      l.channels.push_back(h);
    }
    return l;
  }

  ExtraSpeakers extraPosVerticalNominal(const HostLayout &layout) {
    ExtraSpeakers out;
    const int S = (int)layout.channels.size();
    for (int i = 0; i < S; i++) out.parent.push_back(i);

    std::vector<int> mid;
    for (int i = 0; i < S; i++) {
      double el = layout.channels[i].nominal.el;
      if (-10.0 <= el && el <= 10.0) mid.push_back(i);
    }
    const double layers[2][3] = {{-30.0, -70.0, -10.0}, {30.0, 10.0, 70.0}};
    for (const auto &layer : layers) {
      const double nominal_el = layer[0], lo = layer[1], hi = layer[2];
      std::vector<int> cur;
      for (int i = 0; i < S; i++) {
        double el = layout.channels[i].nominal.el;
        if (lo <= el && el <= hi) cur.push_back(i);
      }
      // mid-layer speakers further round than the layer's azimuth range (+40 deg) get
      // a copy at this layer's elevation, downmixed 1:1 into them; an empty layer
      // copies every mid-layer speaker.
      double az_limit = 0.0, real_el = nominal_el;
      if (!cur.empty()) {
        double az_range = std::numeric_limits<double>::min();
        double el_sum = 0.0;
        for (int i : cur) {
          az_range = std::max(az_range, std::abs(layout.channels[i].nominal.az));
          el_sum += layout.channels[i].pos.el;
        }
        az_limit = az_range + 40.0;
        real_el = el_sum / (double)cur.size();
      }
      const double eps = 1e-5;
      for (int i : mid) {
        const HostChannel &m = layout.channels[i];
        if (std::abs(m.pos.az) >= az_limit - eps) {
          HostChannel e;
          e.name = "extra";
          e.pos = {m.pos.az, real_el, 1.0};
          e.nominal = {m.nominal.az, nominal_el, 1.0};
          e.az_lo = e.az_hi = e.nominal.az;
          e.el_lo = e.el_hi = e.nominal.el;
          e.lfe = false;
          out.channels.push_back(e);
          // the reference looks the parent up by *name* (first match)
          out.parent.push_back(layout.indexOf(m.name));
        }
      }
    }
    return out;
  }

  namespace {
    // Sum of one column of an n x 3 column-major matrix in the order Eigen's SSE2
    // (2-double packet) linear-vectorised reduction uses -- the libear reference takes
    // the centre as vertices.colwise().mean() (ref geom.cpp:39).  Mirror-symmetric
    // layouts make the "least collinear vertex" choice below an exact tie that only this
    // rounding breaks, and the choice fixes the scan order of a virtual n-gon's inner
    // triangles, so the order is reproduced rather than approximated.  `col` decides the
    // alignment of the column inside the 16-byte aligned matrix: elements before the
    // first aligned one and after the last full packet are added last, one by one.
    double eigenColumnSum(const double *v, int n, int col) {
      const int start = (col * n) & 1;
      const int aligned2 = ((n - start) / 4) * 4, aligned = ((n - start) / 2) * 2;
      const int end2 = start + aligned2, end = start + aligned;
      if (!aligned) {
        double r = v[0];
        for (int i = 1; i < n; i++) r += v[i];
        return r;
      }
      double p0[2] = {v[start], v[start + 1]};
      if (aligned > 2) {
        double p1[2] = {v[start + 2], v[start + 3]};
        for (int i = start + 4; i < end2; i += 4) {
          p0[0] += v[i];
          p0[1] += v[i + 1];
          p1[0] += v[i + 2];
          p1[1] += v[i + 3];
        }
        p0[0] += p1[0];
        p0[1] += p1[1];
        if (end > end2) {
          p0[0] += v[end2];
          p0[1] += v[end2 + 1];
        }
      }
      double r = p0[0] + p0[1];
      for (int i = 0; i < start; i++) r += v[i];
      for (int i = end; i < n; i++) r += v[i];
      return r;
    }
  }  // namespace

  // Facets of the convex hull of a handful of points, brute force (ref src/common/convex_hull.cpp:7-100;
  // same acceptance tests and tolerance, so the same facets in the same order): every triangle
  // (i < j < k) whose plane has all points on the side of the centroid is a hull triangle;
  // coplanar hull triangles merge into one facet (sorted vertex set) in order of first appearance.
  std::vector<std::vector<int>> convexHull(const std::vector<Vec3> &pts, double tol) {
    const int n = (int)pts.size();
    if (n < 4) throw Error(INVALID_ARGUMENT, "convex hull needs at least 4 loudspeaker positions");
    Vec3 inside = {0.0, 0.0, 0.0};
    for (const Vec3 &p : pts) inside = inside + p;
    inside = {inside.x / (double)n, inside.y / (double)n, inside.z / (double)n};

    struct HullFacet {
      std::set<int> verts;
      Vec3 normal;
    };
    std::vector<HullFacet> hull;
    for (int i = 0; i < n; i++)
      for (int j = i + 1; j < n; j++)
        for (int k = j + 1; k < n; k++) {
          Vec3 nrm = cross(pts[j] - pts[i], pts[k] - pts[i]);
          const double sq = dot(nrm, nrm);
          if (!(sq > tol)) throw Error(INTERNAL_ERROR, "collinear points in convex hull");
          const double len = std::sqrt(sq);
          nrm = {nrm.x / len, nrm.y / len, nrm.z / len};
          const double d_inside = dot(nrm, inside - pts[i]);
          if (std::abs(d_inside) < tol) continue;  // plane through the centroid: not a hull face
          bool all_inside = true;
          for (const Vec3 &q : pts) {
            const double d = dot(nrm, q - pts[i]);
            if (!(d_inside > 0 ? d > -tol : d < tol)) {
              all_inside = false;
              break;
            }
          }
          if (!all_inside) continue;
          // same plane as an earlier facet? (its lowest vertex lies in this plane and the
          // normals are parallel)
          bool merged = false;
          for (HullFacet &f : hull) {
            const Vec3 &fp = pts[*f.verts.begin()];
            const Vec3 c = cross(f.normal, nrm);
            if (std::abs(dot(pts[i] - fp, f.normal)) < tol && dot(c, c) < tol) {
              f.verts.insert({i, j, k});
              merged = true;
              break;
            }
          }
          if (!merged) hull.push_back({{i, j, k}, nrm});
        }
    std::vector<std::vector<int>> out;
    for (const HullFacet &f : hull) out.emplace_back(f.verts.begin(), f.verts.end());
    return out;
  }

  std::vector<int> ngonVertexOrder(const std::vector<Vec3> &verts) {
    const int n = (int)verts.size();
    std::vector<double> col(n);
    double mean[3];
    for (int k = 0; k < 3; k++) {
      for (int i = 0; i < n; i++) col[i] = k == 0 ? verts[i].x : (k == 1 ? verts[i].y : verts[i].z);
      mean[k] = eigenColumnSum(col.data(), n, k) / (double)n;
    }
    const Vec3 centre = {mean[0], mean[1], mean[2]};
    // two in-plane axes: the first vertex, and the vertex least collinear with it
    Vec3 a = verts[0] - centre;
    Vec3 b = {0, 0, 0};
    double best = std::numeric_limits<double>::max();
    for (int i = 1; i < n; i++) {
      Vec3 v = verts[i] - centre;
      double c = std::abs(dot(v, a));
      if (c < best) {
        best = c;
        b = v;
      }
    }
    std::vector<double> angle(n);
    for (int i = 0; i < n; i++) {
      Vec3 v = verts[i] - centre;
      angle[i] = std::atan2(dot(v, a), dot(v, b));
    }
    std::vector<int> order(n);
    for (int i = 0; i < n; i++) order[i] = i;
    std::sort(order.begin(), order.end(), [&](int i, int j) { return angle[i] < angle[j]; });
    return order;
  }

  void buildPspTables(const HostLayout &layout_in, PspTables &out) {
    HostLayout layout = layout_in.withoutLfe();
    checkScreenSpeakers(layout);
    if (layout.name == "0+2+0") {
      // stereo pans on 0+5+0 and downmixes (ref point_source_panner.cpp:362-392, 400-423)
      if (layout.indexOf("M+030") != 0 || layout.indexOf("M-030") != 1 ||
          layout.channels.size() != 2)
        throw Error(INVALID_ARGUMENT, "0+2+0 must consist of M+030, M-030");
      buildFull(getLayout("0+5+0").withoutLfe(), out);
      out.stereo = 1;
      out.n_real = 2;
    } else {
      buildFull(layout, out);
    }
  }

  std::vector<Vec3> extentGridPositions(int n_rows) {
    std::vector<Vec3> out;
    for (int r = 0; r < n_rows; r++) {
      // Eigen LinSpaced(n, -90, 90): low + i*step, last element exactly high
      double step = 180.0 / (double)(n_rows - 1);
      double el = (r == n_rows - 1) ? 90.0 : -90.0 + (double)r * step;
      double radius = std::cos(radians(el));
      double perimiter = 2 * kPi * radius;
      double perimiter_centre = 2 * kPi;
      int n_points = (int)std::round((perimiter / perimiter_centre) * 2 * (n_rows - 1));
      if (n_points == 0) n_points = 1;
      double az_step = 360.0 / (double)n_points;
      for (int i = 0; i < n_points; i++) out.push_back(cart((double)i * az_step, el, 1.0));
    }
    return out;
  }

}  // namespace earb

namespace earb {
  namespace {
    const double kTDesign[][2] = {
#include "tdesign_5200.inc"
    };
  }  // namespace

  const std::vector<Vec3> &hoaDesignPoints() {
    static const std::vector<Vec3> pts = [] {
      std::vector<Vec3> p;
      const size_t n = sizeof(kTDesign) / sizeof(kTDesign[0]);
      for (size_t i = 0; i < n; i++) {
        double phi = kTDesign[i][0], theta = kTDesign[i][1];
        p.push_back({std::sin(theta) * std::cos(phi), std::sin(theta) * std::sin(phi),
                     std::cos(theta)});
      }
      return p;
    }();
    return pts;
  }
}  // namespace earb

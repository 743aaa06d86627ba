// TEST INFRASTRUCTURE ONLY -- never linked into or called from the product path.
//
// extern "C" wrapper over the UNMODIFIED libear reference (compiled from
// /root/reference by oracle/Makefile into oracle/_ref/libear_ref.so) so that the
// Python tests / the cpu_baseline leg of bench.py can drive the reference through
// ctypes.  Every entry point calls the reference's own classes; nothing here
// re-implements arithmetic.  The only addition is the *region index* probe
// (ref_psp_handle): libear does not expose which region handled a position
// (PolarPointSourcePanner::_regions is private, point_source_panner.hpp:143), so
// the region list is rebuilt exactly as configureFullPolarPanner does
// (point_source_panner.cpp:473-557) from the reference's own linkable pieces
// (getAugmentedLayout, FACETS, Triplet, QuadRegion, VirtualNgon) and the first
// region whose handleRemap() returns a value is reported.
#include <complex>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

// everything point_source_panner.hpp includes is pulled in first (all #pragma
// once), so the access override below only touches that header's own classes
#include <Eigen/Core>
#include <numeric>
#include <set>
#include "ear/layout.hpp"
#include "ear/metadata.hpp"
#define private public  // VirtualNgon::_regions (inner triplet index probe)
#include "common/point_source_panner.hpp"
#undef private

#include "common/convex_hull.hpp"
#include "common/facets.hpp"
#include "common/geom.hpp"
#include "common/helpers/eigen_helpers.hpp"
#include "direct_speakers/gain_calculator_direct_speakers.hpp"
#include "direct_speakers/mapping_rules.hpp"
#include "ear/bs2051.hpp"
#include "ear/conversion.hpp"
#include "ear/decorrelate.hpp"
#include "ear/dsp/block_convolver.hpp"
#include "ear/dsp/delay_buffer.hpp"
#include "ear/dsp/gain_interpolator.hpp"
#include "ear/dsp/variable_block_size.hpp"
#include "ear/fft.hpp"
#include "ear/gain_calculators.hpp"
#include "hoa/hoa.hpp"
#include "object_based/polar_extent.hpp"
#include "reference/extent.hpp"

using namespace ear;

namespace ear {
  Facet _adjacent_verts(std::vector<Facet> facets, int vert);  // point_source_panner.cpp:350
}

static thread_local std::string g_err;

#define REF_TRY try {
#define REF_CATCH                                   \
  }                                                 \
  catch (const ear::not_implemented &e) {           \
    g_err = e.what();                               \
    return 1;                                       \
  }                                                 \
  catch (const ear::internal_error &e) {            \
    g_err = e.what();                               \
    return 2;                                       \
  }                                                 \
  catch (const ear::adm_error &e) {                 \
    g_err = e.what();                               \
    return 3;                                       \
  }                                                 \
  catch (const ear::unknown_layout &e) {            \
    g_err = e.what();                               \
    return 4;                                       \
  }                                                 \
  catch (const ear::invalid_argument &e) {          \
    g_err = e.what();                               \
    return 5;                                       \
  }                                                 \
  catch (const std::exception &e) {                 \
    g_err = e.what();                               \
    return 6;                                       \
  }

struct RefPsp {
  std::shared_ptr<PointSourcePanner> psp;
  std::vector<std::unique_ptr<RegionHandler>> regions;  // rebuilt list
  int n_aug = 0;
  Eigen::MatrixXd downmix;
};

struct RefExtent {
  std::shared_ptr<PointSourcePanner> psp;
  std::unique_ptr<PolarExtent> ext;
  std::unique_ptr<reference::PolarExtentPanner> ext_double;
};

struct RefConvolver {
  std::unique_ptr<dsp::block_convolver::Context> ctx;
  std::vector<std::unique_ptr<dsp::block_convolver::Filter>> filters;
  std::unique_ptr<dsp::block_convolver::BlockConvolver> conv;
  size_t block_size;
};

extern "C" {

const char *ref_last_error() { return g_err.c_str(); }

// ---- layouts -----------------------------------------------------------
int ref_num_layouts() { return (int)loadLayouts().size(); }
int ref_layout_name(int i, char *buf, int buflen) {
  std::strncpy(buf, loadLayouts().at(i).name().c_str(), buflen);
  return 0;
}
// names: C strings of 16 bytes each; pos/nominal: [C][3] az,el,dist; ranges [C][4]
int ref_layout_info(const char *name, int *n_channels, int *is_lfe, double *pos,
                    double *nominal, double *ranges, char *names) {
  REF_TRY
  Layout l = getLayout(name);
  *n_channels = (int)l.channels().size();
  if (!is_lfe) return 0;
  for (size_t i = 0; i < l.channels().size(); i++) {
    const Channel &c = l.channels()[i];
    is_lfe[i] = c.isLfe();
    pos[3 * i] = c.polarPosition().azimuth;
    pos[3 * i + 1] = c.polarPosition().elevation;
    pos[3 * i + 2] = c.polarPosition().distance;
    nominal[3 * i] = c.polarPositionNominal().azimuth;
    nominal[3 * i + 1] = c.polarPositionNominal().elevation;
    nominal[3 * i + 2] = c.polarPositionNominal().distance;
    ranges[4 * i] = c.azimuthRange().first;
    ranges[4 * i + 1] = c.azimuthRange().second;
    ranges[4 * i + 2] = c.elevationRange().first;
    ranges[4 * i + 3] = c.elevationRange().second;
    std::strncpy(names + 16 * i, c.name().c_str(), 16);
  }
  return 0;
  REF_CATCH
}

// facets of a layout in FACETS order, each facet's vertices in std::set (ascending)
// order; out: flat int array, each facet = [count, v0, v1, ...]; returns length via *n
int ref_facets(const char *layout_name, int *out, int cap, int *n) {
  REF_TRY
  const auto &f = FACETS.at(layout_name);
  int k = 0;
  for (const auto &facet : f) {
    if (k + 1 + (int)facet.size() > cap) return 8;
    out[k++] = (int)facet.size();
    for (int v : facet) out[k++] = v;
  }
  *n = k;
  return 0;
  REF_CATCH
}

// ---- point source panner + region index --------------------------------
// a Layout from flat arrays: names separated by '\n'; pos / nominal [n][3]; ranges [n][2]
static Layout make_custom_layout(const char *name, int n, const char *names, const double *pos,
                                 const double *nominal, const double *az_range,
                                 const double *el_range, const int *is_lfe) {
  std::vector<Channel> channels;
  std::string all(names);
  size_t start = 0;
  for (int i = 0; i < n; i++) {
    size_t end = all.find('\n', start);
    std::string cname = all.substr(start, end == std::string::npos ? std::string::npos : end - start);
    start = end == std::string::npos ? all.size() : end + 1;
    channels.emplace_back(
        cname, PolarPosition{pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]},
        PolarPosition{nominal[3 * i], nominal[3 * i + 1], nominal[3 * i + 2]},
        std::make_pair(az_range[2 * i], az_range[2 * i + 1]),
        std::make_pair(el_range[2 * i], el_range[2 * i + 1]), is_lfe[i] != 0);
  }
  return Layout(name, channels);
}

static int build_psp(Layout layout, void **out);

int ref_psp_new(const char *layout_name, void **out) {
  REF_TRY
  return build_psp(getLayout(layout_name).withoutLfe(), out);
  REF_CATCH
}

// custom layout (any name): the hull comes from convex_hull when the name has no FACETS entry
int ref_psp_new_custom(const char *name, int n, const char *names, const double *pos,
                       const double *nominal, const double *az_range, const double *el_range,
                       const int *is_lfe, void **out) {
  REF_TRY
  return build_psp(make_custom_layout(name, n, names, pos, nominal, az_range, el_range, is_lfe).withoutLfe(), out);
  REF_CATCH
}

// ear::convex_hull on n points; out: flat int array, each facet = [count, v0, v1, ...]
int ref_convex_hull(int n, const double *xyz, int *out, int cap, int *n_out) {
  REF_TRY
  std::vector<Eigen::Vector3d> pts;
  for (int i = 0; i < n; i++) pts.emplace_back(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
  int k = 0;
  for (const auto &facet : convex_hull(pts)) {
    if (k + 1 + (int)facet.size() > cap) return 8;
    out[k++] = (int)facet.size();
    for (int v : facet) out[k++] = v;
  }
  *n_out = k;
  return 0;
  REF_CATCH
}

static int build_psp(Layout layout, void **out) {
  {
  auto h = std::make_unique<RefPsp>();
  h->psp = configurePolarPanner(layout);
  if (layout.name() != "0+2+0") {
    std::vector<Eigen::Vector3d> positionsReal, positionsNominal;
    std::set<int> virtualVerts;
    std::tie(positionsReal, positionsNominal, virtualVerts, h->downmix) =
        getAugmentedLayout(layout);
    auto facets_it = FACETS.find(layout.name());
    std::vector<Facet> facets =
        facets_it != FACETS.end() ? facets_it->second : convex_hull(positionsNominal);
    for (int virtualVert : virtualVerts) {
      Facet realVerts = _adjacent_verts(facets, virtualVert);
      std::vector<int> realVertsVec(realVerts.begin(), realVerts.end());
      Eigen::Map<Eigen::VectorXi> outputChannels(realVertsVec.data(),
                                                 realVertsVec.size());
      Eigen::MatrixXd positions(outputChannels.size(), 3);
      int rowIndex = 0;
      for (int vert : realVerts) positions.row(rowIndex++) = positionsReal[vert];
      Eigen::Vector3d centrePosition = positionsReal[virtualVert];
      Eigen::VectorXd centreDownmix(outputChannels.size());
      centreDownmix.fill(1.0 / std::sqrt((double)outputChannels.size()));
      h->regions.push_back(std::make_unique<VirtualNgon>(
          outputChannels, positions, centrePosition, centreDownmix));
    }
    for (const auto &facetVerts : facets) {
      if (doIntersect(facetVerts.begin(), facetVerts.end(), virtualVerts.begin(),
                      virtualVerts.end()))
        continue;
      std::vector<int> v(facetVerts.begin(), facetVerts.end());
      Eigen::MatrixXd positions(v.size(), 3);
      int rowIndex = 0;
      for (int vert : facetVerts) positions.row(rowIndex++) = positionsReal[vert];
      if (v.size() == 3) {
        Eigen::Vector3i oc(v.data());
        h->regions.push_back(std::make_unique<Triplet>(oc, positions));
      } else {
        Eigen::Vector4i oc(v.data());
        h->regions.push_back(std::make_unique<QuadRegion>(oc, positions));
      }
    }
    h->n_aug = (int)h->downmix.rows();
  }
  *out = h.release();
  return 0;
  }
}
void ref_psp_free(void *h) { delete (RefPsp *)h; }
int ref_psp_num_channels(void *h) {
  return ((RefPsp *)h)->psp->numberOfOutputChannels();
}
int ref_psp_num_regions(void *h) { return (int)((RefPsp *)h)->regions.size(); }

// xyz [n][3]; gains [n][S]; region_idx [n] (nullable); inner_idx [n] (nullable):
// inner triplet index for VirtualNgon regions, else -1
int ref_psp_handle(void *hv, long n, const double *xyz, double *gains,
                   int32_t *region_idx, int32_t *inner_idx) {
  REF_TRY
  RefPsp *h = (RefPsp *)hv;
  int S = h->psp->numberOfOutputChannels();
  for (long i = 0; i < n; i++) {
    Eigen::Vector3d pos(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    auto pv = h->psp->handle(pos);
    if (!pv) {
      g_err = "psp returned none";
      return 7;
    }
    for (int s = 0; s < S; s++) gains[i * S + s] = pv.get()(s);
    if (region_idx) {
      int found = -1, inner = -1;
      for (size_t r = 0; r < h->regions.size(); r++) {
        auto rv = h->regions[r]->handleRemap(pos, h->n_aug);
        if (rv) {
          found = (int)r;
          if (auto *ng = dynamic_cast<VirtualNgon *>(h->regions[r].get())) {
            for (size_t k = 0; k < ng->_regions.size(); k++)
              if (ng->_regions[k]->handle(pos)) {
                inner = (int)k;
                break;
              }
          }
          break;
        }
      }
      region_idx[i] = found;
      if (inner_idx) inner_idx[i] = inner;
    }
  }
  return 0;
  REF_CATCH
}

// ---- extent panner -------------------------------------------------------
// core: 0 = default (xsimd dispatch), 1 = scalar float core, 2 = the double
// precision restatement in the reference's tests (tests/reference/extent.cpp)
int ref_extent_new(const char *layout_name, int core, void **out) {
  REF_TRY
  Layout layout = getLayout(layout_name).withoutLfe();
  auto h = std::make_unique<RefExtent>();
  h->psp = configurePolarPanner(layout);
  if (core == 0)
    h->ext = std::make_unique<PolarExtent>(h->psp);
  else if (core == 1)
    h->ext = std::make_unique<PolarExtent>(h->psp, get_polar_extent_core_scalar());
  else
    h->ext_double = std::make_unique<reference::PolarExtentPanner>(h->psp);
  *out = h.release();
  return 0;
  REF_CATCH
}
void ref_extent_free(void *h) { delete (RefExtent *)h; }
int ref_extent_handle(void *hv, long n, const double *xyz, const double *width,
                      const double *height, const double *depth, double *out) {
  REF_TRY
  RefExtent *h = (RefExtent *)hv;
  int S = h->psp->numberOfOutputChannels();
  Eigen::VectorXd pv(S);
  for (long i = 0; i < n; i++) {
    Eigen::Vector3d pos(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    if (h->ext)
      h->ext->handle(pos, width[i], height[i], depth[i], pv);
    else
      pv = h->ext_double->handle(pos, width[i], height[i], depth[i]);
    for (int s = 0; s < S; s++) out[i * S + s] = pv(s);
  }
  return 0;
  REF_CATCH
}
// name of the ISA xsimd::dispatch selects on this host (informational)
const char *ref_extent_isa() {
  auto core = get_polar_extent_core();
  switch (core->batch_size()) {
    case 16: return "avx512bw";
    case 8: return "avx/avx2";
    case 4: return "sse";
    case 1: return "scalar";
    default: return "unknown";
  }
}

// ---- GainCalculatorObjects ---------------------------------------------
int ref_objects_new(const char *layout_name, void **out) {
  REF_TRY
  *out = new GainCalculatorObjects(getLayout(layout_name));
  return 0;
  REF_CATCH
}
int ref_objects_new_custom(const char *name, int n, const char *names, const double *pos,
                           const double *nominal, const double *az_range, const double *el_range,
                           const int *is_lfe, void **out) {
  REF_TRY
  *out = new GainCalculatorObjects(make_custom_layout(name, n, names, pos, nominal, az_range, el_range, is_lfe));
  return 0;
  REF_CATCH
}
void ref_objects_free(void *h) { delete (GainCalculatorObjects *)h; }

// flags bit0: cartesian flag, bit1: channelLock, bit2: screenRef, bit3: divergence!=0,
// bit4: zoneExclusion non-empty, bit5: position is Cartesian variant
static ObjectsTypeMetadata make_otm(double az, double el, double dist, double w,
                                    double h, double d, double gain,
                                    double diffuse, int flags) {
  ObjectsTypeMetadata otm;
  if (flags & 32)
    otm.position = CartesianPosition(0.0, 1.0, 0.0);
  else
    otm.position = PolarPosition(az, el, dist);
  otm.width = w;
  otm.height = h;
  otm.depth = d;
  otm.gain = gain;
  otm.diffuse = diffuse;
  if (flags & 1) otm.cartesian = true;
  if (flags & 2) otm.channelLock.flag = true;
  if (flags & 4) otm.screenRef = true;
  if (flags & 8) otm.objectDivergence = PolarObjectDivergence(0.5);
  if (flags & 16) otm.zoneExclusion.zones.push_back(PolarExclusionZone{});
  return otm;
}

// SoA inputs, [n][C_out] row-major outputs. is_float selects calculate<float>.
int ref_objects_calc(void *hv, long n, const double *az, const double *el,
                     const double *dist, const double *width,
                     const double *height, const double *depth,
                     const double *gain, const double *diffuse, const int *flags,
                     int n_out, int is_float, void *direct, void *diffuse_out) {
  REF_TRY
  auto *gc = (GainCalculatorObjects *)hv;
  if (is_float) {
    std::vector<float> d(n_out), f(n_out);
    for (long i = 0; i < n; i++) {
      auto otm = make_otm(az[i], el[i], dist[i], width[i], height[i], depth[i],
                          gain[i], diffuse[i], flags ? flags[i] : 0);
      gc->calculate(otm, d, f);
      std::memcpy((float *)direct + i * n_out, d.data(), n_out * sizeof(float));
      std::memcpy((float *)diffuse_out + i * n_out, f.data(), n_out * sizeof(float));
    }
  } else {
    std::vector<double> d(n_out), f(n_out);
    for (long i = 0; i < n; i++) {
      auto otm = make_otm(az[i], el[i], dist[i], width[i], height[i], depth[i],
                          gain[i], diffuse[i], flags ? flags[i] : 0);
      gc->calculate(otm, d, f);
      std::memcpy((double *)direct + i * n_out, d.data(), n_out * sizeof(double));
      std::memcpy((double *)diffuse_out + i * n_out, f.data(), n_out * sizeof(double));
    }
  }
  return 0;
  REF_CATCH
}

// ---- GainCalculatorHOA ----------------------------------------------------
// D: [C][C_out] (one vector of per-loudspeaker gains per input channel)
int ref_hoa_calc(const char *layout_name, int C, const int *orders,
                 const int *degrees, const char *norm, double nfc_ref_dist,
                 int screen_ref, int n_out, double *D, int *n_warnings) {
  REF_TRY
  GainCalculatorHOA gc(getLayout(layout_name));
  HOATypeMetadata md;
  md.orders.assign(orders, orders + C);
  md.degrees.assign(degrees, degrees + C);
  md.normalization = norm;
  md.nfcRefDist = nfc_ref_dist;
  md.screenRef = screen_ref != 0;
  std::vector<std::vector<double>> gains(C, std::vector<double>(n_out));
  int nw = 0;
  gc.calculate(md, gains, [&](const Warning &) { nw++; });
  for (int c = 0; c < C; c++)
    for (int o = 0; o < n_out; o++) D[c * n_out + o] = gains[c][o];
  if (n_warnings) *n_warnings = nw;
  return 0;
  REF_CATCH
}
int ref_hoa_num_points() { return (int)hoa::load_points().rows(); }
int ref_hoa_points(double *xyz) {
  auto p = hoa::load_points();
  for (Eigen::Index i = 0; i < p.rows(); i++)
    for (int k = 0; k < 3; k++) xyz[3 * i + k] = p(i, k);
  return 0;
}

// ---- GainCalculatorDirectSpeakers ----------------------------------------
// labels: '\n'-separated list. position polar (az, el, dist) with optional bounds
// (bounds[6]: azMin, azMax, elMin, elMax, distMin, distMax; has_bounds bitmask).
int ref_direct_speakers_calc(const char *layout_name, const char *labels,
                             double az, double el, double dist,
                             const double *bounds, int has_bounds,
                             double low_pass, double high_pass, int has_freq,
                             const char *pack_id, int n_out, double *gains,
                             int *n_warnings, const char *substitutions,
                             int cartesian, int edge_lock) {
  REF_TRY
  // additionalSubstitutions: '\n'-separated "from=to" pairs
  std::map<std::string, std::string> subs;
  std::string ss(substitutions ? substitutions : "");
  for (size_t b = 0; b < ss.size();) {
    size_t e = ss.find('\n', b);
    if (e == std::string::npos) e = ss.size();
    std::string kv = ss.substr(b, e - b);
    size_t eq = kv.find('=');
    if (eq != std::string::npos) subs[kv.substr(0, eq)] = kv.substr(eq + 1);
    b = e + 1;
  }
  GainCalculatorDirectSpeakers gc(getLayout(layout_name), subs);
  DirectSpeakersTypeMetadata md;
  std::string ls(labels ? labels : "");
  size_t start = 0;
  while (start < ls.size()) {
    size_t e = ls.find('\n', start);
    if (e == std::string::npos) e = ls.size();
    md.speakerLabels.push_back(ls.substr(start, e - start));
    start = e + 1;
  }
  PolarSpeakerPosition p(az, el, dist);
  if (has_bounds & 1) p.azimuthMin = boost::optional<double>(bounds[0]);
  if (has_bounds & 2) p.azimuthMax = boost::optional<double>(bounds[1]);
  if (has_bounds & 4) p.elevationMin = boost::optional<double>(bounds[2]);
  if (has_bounds & 8) p.elevationMax = boost::optional<double>(bounds[3]);
  if (has_bounds & 16) p.distanceMin = boost::optional<double>(bounds[4]);
  if (has_bounds & 32) p.distanceMax = boost::optional<double>(bounds[5]);
  if (edge_lock & 1) p.screenEdgeLock.horizontal = std::string("left");
  if (edge_lock & 2) p.screenEdgeLock.vertical = std::string("top");
  md.position = p;
  if (cartesian) md.position = CartesianSpeakerPosition(az, el, dist);
  if (has_freq & 1) md.channelFrequency.lowPass = boost::optional<double>(low_pass);
  if (has_freq & 2) md.channelFrequency.highPass = boost::optional<double>(high_pass);
  if (pack_id && pack_id[0]) md.audioPackFormatID = boost::optional<std::string>(std::string(pack_id));
  std::vector<double> g(n_out);
  int nw = 0;
  // n_warnings: count in the low byte, bitmask of (1 << code) above it
  gc.calculate(md, g, [&](const Warning &w) { nw = (nw + 1) | (256 << (int)w.code); });
  for (int o = 0; o < n_out; o++) gains[o] = g[o];
  if (n_warnings) *n_warnings = nw;
  return 0;
  REF_CATCH
}

// mapping rules / ITU pack table as text (data export for tools/gen_mapping_rules.py):
// one rule per line "R|label|ch=gain,ch=gain|in,in|out,out", one pack per line "P|id|layout"
int ref_mapping_rules_dump(char *buf, long cap, long *needed) {
  REF_TRY
  std::string s;
  char num[64];
  for (const MappingRule &r : rules) {
    s += "R|" + r.speakerLabel + "|";
    for (size_t i = 0; i < r.gains.size(); i++) {
      snprintf(num, sizeof num, "%.17g", r.gains[i].second);
      s += (i ? "," : "") + r.gains[i].first + "=" + num;
    }
    s += "|";
    for (size_t i = 0; i < r.input_layouts.size(); i++) s += (i ? "," : "") + r.input_layouts[i];
    s += "|";
    for (size_t i = 0; i < r.output_layouts.size(); i++) s += (i ? "," : "") + r.output_layouts[i];
    s += "\n";
  }
  for (const auto &kv : itu_packs) s += "P|" + kv.first + "|" + kv.second + "\n";
  if (needed) *needed = (long)s.size() + 1;
  if (buf && cap > (long)s.size()) memcpy(buf, s.c_str(), s.size() + 1);
  return 0;
  REF_CATCH
}

// ---- decorrelators --------------------------------------------------------
int ref_decorrelators(const char *layout_name, int n_channels, float *out_f,
                      double *out_d) {
  REF_TRY
  Layout l = getLayout(layout_name);
  if (out_f) {
    auto f = designDecorrelators<float>(l);
    for (int c = 0; c < n_channels; c++)
      std::memcpy(out_f + 512 * c, f.at(c).data(), 512 * sizeof(float));
  }
  if (out_d) {
    auto f = designDecorrelators<double>(l);
    for (int c = 0; c < n_channels; c++)
      std::memcpy(out_d + 512 * c, f.at(c).data(), 512 * sizeof(double));
  }
  return 0;
  REF_CATCH
}
int ref_decorrelator_delay() { return decorrelatorCompensationDelay(); }

// ---- GainInterpolator ------------------------------------------------------
// in: [N][T] planar, out: [M][T] planar. points: sample index + [N][M] gains each.
// The buffer is pushed through process() in calls of `call_size` samples, starting
// at sample index block_start0.
int ref_interp_matrix(int N, int M, long T, const float *in, float *out,
                      int n_points, const long *point_sample,
                      const float *point_gains, long block_start0,
                      long call_size) {
  REF_TRY
  dsp::GainInterpolator<dsp::LinearInterpMatrix> interp;
  for (int p = 0; p < n_points; p++) {
    std::vector<std::vector<float>> pt(N, std::vector<float>(M));
    for (int n = 0; n < N; n++)
      for (int m = 0; m < M; m++)
        pt[n][m] = point_gains[((size_t)p * N + n) * M + m];
    interp.interp_points.emplace_back(point_sample[p], std::move(pt));
  }
  std::vector<const float *> in_p(N);
  std::vector<float *> out_p(M);
  for (long t = 0; t < T; t += call_size) {
    long n = std::min(call_size, T - t);
    for (int c = 0; c < N; c++) in_p[c] = in + (size_t)c * T + t;
    for (int c = 0; c < M; c++) out_p[c] = out + (size_t)c * T + t;
    interp.process(block_start0 + t, (size_t)n, in_p.data(), out_p.data());
  }
  return 0;
  REF_CATCH
}
// LinearInterpVector: in [T], out [M][T]; point gains [n_points][M]
int ref_interp_vector(int M, long T, const float *in, float *out, int n_points,
                      const long *point_sample, const float *point_gains,
                      long block_start0, long call_size) {
  REF_TRY
  dsp::GainInterpolator<dsp::LinearInterpVector> interp;
  for (int p = 0; p < n_points; p++)
    interp.interp_points.emplace_back(
        point_sample[p],
        std::vector<float>(point_gains + (size_t)p * M,
                           point_gains + (size_t)(p + 1) * M));
  std::vector<float *> out_p(M);
  for (long t = 0; t < T; t += call_size) {
    long n = std::min(call_size, T - t);
    const float *in_p = in + t;
    for (int c = 0; c < M; c++) out_p[c] = out + (size_t)c * T + t;
    interp.process(block_start0 + t, (size_t)n, &in_p, out_p.data());
  }
  return 0;
  REF_CATCH
}
// LinearInterpSingle: in [T], out [T]; point gains [n_points]
int ref_interp_single(long T, const float *in, float *out, int n_points,
                      const long *point_sample, const float *point_gains,
                      long block_start0, long call_size) {
  REF_TRY
  dsp::GainInterpolator<dsp::LinearInterpSingle> interp;
  for (int p = 0; p < n_points; p++)
    interp.interp_points.emplace_back(point_sample[p], point_gains[p]);
  for (long t = 0; t < T; t += call_size) {
    long n = std::min(call_size, T - t);
    const float *in_p = in + t;
    float *out_p = out + t;
    interp.process(block_start0 + t, (size_t)n, &in_p, &out_p);
  }
  return 0;
  REF_CATCH
}

// ---- BlockConvolver -----------------------------------------------------
int ref_conv_new(long block_size, long num_blocks, void **out) {
  REF_TRY
  auto h = std::make_unique<RefConvolver>();
  h->block_size = block_size;
  h->ctx = std::make_unique<dsp::block_convolver::Context>(
      block_size, get_fft_kiss<float>());
  h->conv = std::make_unique<dsp::block_convolver::BlockConvolver>(*h->ctx,
                                                                   num_blocks);
  *out = h.release();
  return 0;
  REF_CATCH
}
void ref_conv_free(void *h) { delete (RefConvolver *)h; }
// returns filter id (>=0) via *id
int ref_conv_add_filter(void *hv, long n, const float *taps, int *id,
                        int *num_blocks) {
  REF_TRY
  auto *h = (RefConvolver *)hv;
  h->filters.push_back(
      std::make_unique<dsp::block_convolver::Filter>(*h->ctx, n, taps));
  *id = (int)h->filters.size() - 1;
  if (num_blocks) *num_blocks = (int)h->filters.back()->num_blocks();
  return 0;
  REF_CATCH
}
// filter_id < 0 means "no filter"
int ref_conv_set_filter(void *hv, int filter_id) {
  REF_TRY
  auto *h = (RefConvolver *)hv;
  if (filter_id < 0)
    h->conv->unset_filter();
  else
    h->conv->set_filter(*h->filters.at(filter_id));
  return 0;
  REF_CATCH
}
int ref_conv_crossfade_filter(void *hv, int filter_id) {
  REF_TRY
  auto *h = (RefConvolver *)hv;
  if (filter_id < 0)
    h->conv->fade_down();
  else
    h->conv->crossfade_filter(*h->filters.at(filter_id));
  return 0;
  REF_CATCH
}
// process n_blocks consecutive blocks; in may be NULL (silence)
int ref_conv_process(void *hv, long n_blocks, const float *in, float *out) {
  REF_TRY
  auto *h = (RefConvolver *)hv;
  for (long b = 0; b < n_blocks; b++)
    h->conv->process(in ? in + b * h->block_size : nullptr,
                     out + b * h->block_size);
  return 0;
  REF_CATCH
}

// ---- DelayBuffer ----------------------------------------------------------
int ref_delay_process(int n_channels, long delay, long T, long call_size,
                      const float *in, float *out) {
  REF_TRY
  dsp::DelayBuffer db(n_channels, delay);
  std::vector<const float *> in_p(n_channels);
  std::vector<float *> out_p(n_channels);
  for (long t = 0; t < T; t += call_size) {
    long n = std::min(call_size, T - t);
    for (int c = 0; c < n_channels; c++) {
      in_p[c] = in + (size_t)c * T + t;
      out_p[c] = out + (size_t)c * T + t;
    }
    db.process(n, in_p.data(), out_p.data());
  }
  return 0;
  REF_CATCH
}

// ---- ear::conversion ------------------------------------------------------
// pos / ext [n][3]; ext nullable (point conversion only); to_polar: Cartesian -> polar
int ref_conversion(long n, int to_polar, const double *pos, const double *ext,
                   double *pos_out, double *ext_out) {
  REF_TRY
  for (long i = 0; i < n; i++) {
    const double *p = pos + 3 * i;
    if (to_polar) {
      CartesianPosition c{p[0], p[1], p[2]};
      PolarPosition r;
      if (ext) {
        conversion::ExtentParams e{ext[3 * i], ext[3 * i + 1], ext[3 * i + 2]}, eo;
        std::tie(r, eo) = conversion::extentCartToPolar(c, e);
        ext_out[3 * i] = eo.width, ext_out[3 * i + 1] = eo.height, ext_out[3 * i + 2] = eo.depth;
      } else {
        r = conversion::pointCartToPolar(c);
      }
      pos_out[3 * i] = r.azimuth, pos_out[3 * i + 1] = r.elevation, pos_out[3 * i + 2] = r.distance;
    } else {
      PolarPosition c{p[0], p[1], p[2]};
      CartesianPosition r;
      if (ext) {
        conversion::ExtentParams e{ext[3 * i], ext[3 * i + 1], ext[3 * i + 2]}, eo;
        std::tie(r, eo) = conversion::extentPolarToCart(c, e);
        ext_out[3 * i] = eo.width, ext_out[3 * i + 1] = eo.height, ext_out[3 * i + 2] = eo.depth;
      } else {
        r = conversion::pointPolarToCart(c);
      }
      pos_out[3 * i] = r.X, pos_out[3 * i + 1] = r.Y, pos_out[3 * i + 2] = r.Z;
    }
  }
  return 0;
  REF_CATCH
}

// ---- FFT plug-in: libear's own KISS transforms (src/fft_kiss.cpp:52-99) -----
// forward: batch x n_fft reals -> batch x (n_fft/2+1) complex (interleaved re, im), unpacked bins;
// reverse: unnormalised c2r
int ref_fft_kiss(long n_fft, long batch, int forward, const float *in, float *out) {
  REF_TRY
  auto plan = get_fft_kiss<float>().plan((size_t)n_fft);
  auto work = plan->alloc_workbuf();
  const long nc = n_fft / 2 + 1;
  std::vector<float> re(n_fft);
  std::vector<std::complex<float>> cx(nc);
  for (long b = 0; b < batch; b++) {
    if (forward) {
      std::copy(in + b * n_fft, in + (b + 1) * n_fft, re.begin());
      plan->transform_forward(re.data(), cx.data(), *work);
      for (long k = 0; k < nc; k++) {
        out[2 * (b * nc + k)] = cx[k].real();
        out[2 * (b * nc + k) + 1] = cx[k].imag();
      }
    } else {
      for (long k = 0; k < nc; k++) cx[k] = {in[2 * (b * nc + k)], in[2 * (b * nc + k) + 1]};
      plan->transform_reverse(cx.data(), re.data(), *work);
      std::copy(re.begin(), re.end(), out + b * n_fft);
    }
  }
  return 0;
  REF_CATCH
}

// ---- VariableBlockSizeAdapter (src/dsp/variable_block_size_impl.cpp:44-81) --
// Runs libear's adapter over `n_calls` calls of the given sizes.  The inner fixed-block process
// function is the probe both sides of the parity test use (tests/cpp/facade_test.cpp, mode "vbs"):
//   out[o][i] = (o + 2) * in[o % n_in][i] + 0.25f * call_index      (call_index counts inner calls)
// in: [n_in][T] planar, out: [n_out][T] planar, T = sum(sizes)
int ref_vbs_process(long block_size, int n_in, int n_out, int n_calls, const long *sizes,
                    const float *in, float *out, int *delay) {
  REF_TRY
  long T = 0;
  for (int c = 0; c < n_calls; c++) T += sizes[c];
  int inner_calls = 0;
  dsp::VariableBlockSizeAdapter adapter(
      (size_t)block_size, (size_t)n_in, (size_t)n_out, [&](const float *const *i, float *const *o) {
        for (int oc = 0; oc < n_out; oc++)
          for (long t = 0; t < block_size; t++)
            o[oc][t] = (float)(oc + 2) * i[oc % n_in][t] + 0.25f * (float)inner_calls;
        inner_calls++;
      });
  if (delay) *delay = adapter.get_delay();
  std::vector<const float *> in_p(n_in);
  std::vector<float *> out_p(n_out);
  long pos = 0;
  for (int c = 0; c < n_calls; c++) {
    for (int k = 0; k < n_in; k++) in_p[k] = in + (size_t)k * T + pos;
    for (int k = 0; k < n_out; k++) out_p[k] = out + (size_t)k * T + pos;
    adapter.process((size_t)sizes[c], in_p.data(), out_p.data());
    pos += sizes[c];
  }
  return 0;
  REF_CATCH
}

}  // extern "C"

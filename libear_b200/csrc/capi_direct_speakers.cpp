// C ABI: DirectSpeakers gain calculator.
//
// Follows ear::GainCalculatorDirectSpeakersImpl::calculate
// (ref src/direct_speakers/gain_calculator_direct_speakers.cpp:244-320): the routing decision
// (mapping rules, label match, closest loudspeaker within bounds, LFE rule) is string and
// table work done on the host; channels that fall through to the point-source panner are
// gathered and panned by K1 (psp_batch_kernel) in one launch per call.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "capi_common.hpp"

using namespace earb;

namespace {
  struct RuleGain {
    const char *channel;
    double gain;
  };
  struct RuleSpec {
    const char *label;
    int gain0, n_gains, in0, n_in, out0, n_out;
  };
  struct PackSpec {
    const char *id, *layout;
  };
#include "mapping_rules.inc"

  // ref src/common/geom.cpp:7-28
  bool insideAngleRange(double x, double start, double end, double tol) {
    while (end - 360.0 > start) end -= 360.0;
    while (end < start) end += 360.0;
    const double start_tol = start - tol;
    while (x - 360.0 >= start_tol) x -= 360.0;
    while (x < start_tol) x += 360.0;
    return x <= end + tol;
  }

  // the capture group of "^urn:itu:bs:2051:[0-9]+:speaker:(.*)$" or the label itself
  // (ref gain_calculator_direct_speakers.hpp:88-89)
  std::string stripUrn(const std::string &label) {
    static const char prefix[] = "urn:itu:bs:2051:";
    static const char middle[] = ":speaker:";
    const size_t np = sizeof prefix - 1, nm = sizeof middle - 1;
    if (label.compare(0, np, prefix) != 0) return label;
    size_t i = np;
    while (i < label.size() && label[i] >= '0' && label[i] <= '9') i++;
    if (i == np || label.compare(i, nm, middle) != 0) return label;
    const std::string rest = label.substr(i + nm);
    if (rest.find_first_of("\n\r") != std::string::npos) return label;  // '.' stops at a line end
    return rest;
  }
}  // namespace

struct earb200_direct_speakers {
  int device = 0;
  HostLayout layout;
  earb200_panner *panner = nullptr;
  int n_out = 0;
  std::vector<Vec3> positions;  // real positions (ref :77)
  std::map<std::string, std::string> substitutions;
  int64_t launches = 0;
  ~earb200_direct_speakers() {
    if (panner) earb200_panner_destroy(panner);
  }

  // ref :138-151
  std::string nominalLabel(const std::string &label) const {
    auto it = substitutions.find(label);
    if (it != substitutions.end()) return it->second;
    return stripUrn(label);
  }

  // ref :26-45
  bool ruleApplies(const RuleSpec &r, const std::string &input_layout, const std::string &label) const {
    if (r.n_in) {
      bool found = false;
      for (int i = 0; i < r.n_in; i++) found |= input_layout == kRuleLayouts[r.in0 + i];
      if (!found) return false;
    }
    if (r.n_out) {
      bool found = false;
      for (int i = 0; i < r.n_out; i++) found |= layout.name == kRuleLayouts[r.out0 + i];
      if (!found) return false;
    }
    if (label != r.label) return false;
    for (int g = 0; g < r.n_gains; g++)
      if (layout.indexOf(kRuleGains[r.gain0 + g].channel) < 0) return false;
    return true;
  }
};

namespace {
  typedef earb200_direct_speakers_metadata Meta;

  enum { WARN_MISMATCH = 1, WARN_NOT_LFE = 2 };  // Warning::Code (ref include/ear/warnings.hpp:13-17)

  // ref :107-136
  bool isLfeChannel(const earb200_direct_speakers &h, const Meta &md, uint32_t &warnings) {
    const bool has_low = md.has_frequency & 1, has_high = md.has_frequency & 2;
    const bool has_lfe_freq = has_low && md.low_pass <= 200 && !has_high;
    if (!has_lfe_freq && (has_low || has_high)) warnings |= 1u << WARN_NOT_LFE;
    bool has_lfe_name = false;
    for (int i = 0; i < md.n_speaker_labels; i++) {
      const std::string nominal = h.nominalLabel(md.speaker_labels[i]);
      if (nominal == "LFE1" || nominal == "LFE2") has_lfe_name = true;
    }
    if (has_lfe_freq != has_lfe_name && md.n_speaker_labels) warnings |= 1u << WARN_MISMATCH;
    return has_lfe_freq || has_lfe_name;
  }

  // ref :153-182 and :222-242: index of the closest loudspeaker of the right type inside the
  // bounds, -1 when there is none or the two closest are within tol of each other
  int channelWithinBounds(const earb200_direct_speakers &h, const Meta &md, bool isLfe, double tol) {
    const double az = md.position[0], el = md.position[1], dist = md.position[2];
    const double azMin = md.has_bounds & 1 ? md.bounds[0] : az, azMax = md.has_bounds & 2 ? md.bounds[1] : az;
    const double elMin = md.has_bounds & 4 ? md.bounds[2] : el, elMax = md.has_bounds & 8 ? md.bounds[3] : el;
    const double dMin = md.has_bounds & 16 ? md.bounds[4] : dist, dMax = md.has_bounds & 32 ? md.bounds[5] : dist;
    const Vec3 p = cart(az, el, dist);
    std::vector<std::pair<int, double>> candidates;
    for (int i = 0; i < h.n_out; i++) {
      const HostChannel &c = h.layout.channels[i];
      if (isLfe != c.lfe) continue;
      if ((insideAngleRange(c.nominal.az, azMin, azMax, tol) || std::abs(c.nominal.el) >= 90.0 - tol) &&
          c.nominal.el > elMin - tol && c.nominal.el < elMax + tol && c.nominal.dist > dMin - tol &&
          c.nominal.dist < dMax + tol) {
        const double dx = h.positions[i].x - p.x, dy = h.positions[i].y - p.y, dz = h.positions[i].z - p.z;
        candidates.push_back(std::make_pair(i, std::sqrt(dx * dx + dy * dy + dz * dz)));
      }
    }
    if (candidates.empty()) return -1;
    if (candidates.size() == 1) return candidates[0].first;
    std::sort(candidates.begin(), candidates.end(),
              [](const std::pair<int, double> &a, const std::pair<int, double> &b) { return a.second < b.second; });
    if (std::abs(candidates[0].second - candidates[1].second) > tol) return candidates[0].first;
    return -1;
  }

  // Routes one channel.  Returns true when the gains are final; false when the channel has to
  // go through the point-source panner at `pan_pos`.
  bool route(const earb200_direct_speakers &h, const Meta &md, double *gains, uint32_t &warnings, Vec3 &pan_pos) {
    if (md.n_speaker_labels < 0 || (md.n_speaker_labels > 0 && !md.speaker_labels))
      throw Error(INVALID_ARGUMENT, "speaker_labels is null");
    if (md.audio_pack_format_id && md.n_speaker_labels == 0)
      throw Error(ADM_ERROR,
                  "ADM error: common definitions audioPackFormatID specified without any speakerLabels as specified in the "
                  "common definitions file");
    if (md.cartesian_position) throw Error(NOT_IMPLEMENTED, "not implemented: Cartesian position");
    const double tol = 1e-5;
    const bool isLfe = isLfeChannel(h, md, warnings);
    for (int c = 0; c < h.n_out; c++) gains[c] = 0.0;

    if (md.audio_pack_format_id) {  // ref :259-276
      const std::string pack = md.audio_pack_format_id;
      for (int k = 0; k < kNumItuPacks; k++) {
        if (pack != kItuPacks[k].id) continue;
        const std::string label = h.nominalLabel(md.speaker_labels[0]);
        for (int r = 0; r < kNumRules; r++) {
          if (!h.ruleApplies(kRules[r], kItuPacks[k].layout, label)) continue;
          for (int g = 0; g < kRules[r].n_gains; g++) {
            const RuleGain &rg = kRuleGains[kRules[r].gain0 + g];
            gains[h.layout.indexOf(rg.channel)] = rg.gain;
          }
          return true;
        }
        break;
      }
    }
    // a loudspeaker with a matching label and type; earlier labels win (ref :278-291)
    for (int i = 0; i < md.n_speaker_labels; i++) {
      const int idx = h.layout.indexOf(h.nominalLabel(md.speaker_labels[i]));
      if (idx >= 0 && isLfe == h.layout.channels[idx].lfe) {
        gains[idx] = 1.0;
        return true;
      }
    }
    // ScreenEdgeLockHandler (ref src/common/screen_edge_lock.hpp:14-20)
    if (md.screen_edge_lock) throw Error(NOT_IMPLEMENTED, "not implemented: screenEdgeLock");
    const int idx = channelWithinBounds(h, md, isLfe, tol);  // ref :295-302
    if (idx >= 0) {
      gains[idx] = 1.0;
      return true;
    }
    if (isLfe) {  // ref :307-313
      const int lfe1 = h.layout.indexOf("LFE1");
      if (lfe1 >= 0) gains[lfe1] = 1.0;
      return true;
    }
    pan_pos = cart(md.position[0], md.position[1], md.position[2]);
    return false;
  }
}  // namespace

extern "C" {

int earb200_direct_speakers_create(const earb200_layout *layout, int device, int n_substitutions,
                                   const char *const *from, const char *const *to,
                                   earb200_direct_speakers **out) {
  return guarded([&] {
    if (!layout || !out) throw Error(INVALID_ARGUMENT, "null argument");
    if (n_substitutions < 0 || (n_substitutions > 0 && (!from || !to)))
      throw Error(INVALID_ARGUMENT, "invalid substitution table");
    // routing is host-side table work; the device is first touched (and must exist) when a
    // channel falls through to the point-source panner
    auto h = std::make_unique<earb200_direct_speakers>();
    h->device = device;
    h->layout = layout->layout;
    h->n_out = (int)h->layout.channels.size();
    for (const HostChannel &c : h->layout.channels) h->positions.push_back(cart(c.pos.az, c.pos.el, c.pos.dist));
    // ref :79-82 (std::map::insert keeps the built-in entries on a key clash)
    h->substitutions = {{"LFE", "LFE1"}, {"LFEL", "LFE1"}, {"LFER", "LFE2"}};
    for (int i = 0; i < n_substitutions; i++) h->substitutions.insert({from[i], to[i]});
    // libear builds the point-source panner in the constructor (configurePolarPanner(layout
    // .withoutLfe()), ref gain_calculator_direct_speakers.cpp:84): layouts it rejects (a screen
    // loudspeaker outside its range, an ill-formed 0+2+0, a degenerate hull) fail HERE, with the
    // same error class.  The region tables are built on the host for that; only their copy to the
    // device waits for the first channel that needs panning.
    {
      PspTables probe;
      buildPspTables(h->layout.withoutLfe(), probe);
    }
    *out = h.release();
  });
}

void earb200_direct_speakers_destroy(earb200_direct_speakers *calc) { delete calc; }

int earb200_direct_speakers_num_outputs(const earb200_direct_speakers *calc) { return calc ? calc->n_out : 0; }

int64_t earb200_direct_speakers_launch_count(const earb200_direct_speakers *calc) {
  return calc ? calc->launches : 0;
}

int earb200_direct_speakers_calculate(earb200_direct_speakers *calc, int64_t n,
                                      const earb200_direct_speakers_metadata *md, double *gains,
                                      uint32_t *warnings) {
  return guarded([&] {
    if (!calc || (n > 0 && (!md || !gains))) throw Error(INVALID_ARGUMENT, "null argument");
    if (n < 0) throw Error(INVALID_ARGUMENT, "negative batch size");
    const int C = calc->n_out;
    std::vector<int64_t> pan_rows;
    std::vector<double> pan_xyz;
    for (int64_t i = 0; i < n; i++) {
      uint32_t w = 0;
      Vec3 p{0, 0, 0};
      const bool done = route(*calc, md[i], gains + i * C, w, p);
      if (warnings) warnings[i] = w;
      if (!done) {
        pan_rows.push_back(i);
        pan_xyz.push_back(p.x);
        pan_xyz.push_back(p.y);
        pan_xyz.push_back(p.z);
      }
    }
    if (pan_rows.empty()) return;
    // point-source panner over the loudspeakers without LFE, scattered past the LFE
    // channels (ref :314-317, mask_write)
    int rc;
    if (!calc->panner) {
      earb200_layout lay{calc->layout};
      rc = earb200_panner_create(&lay, calc->device, &calc->panner);
      if (rc != EARB200_OK) throw Error((Status)rc, last_error());
    }
    const int S = earb200_panner_num_channels(calc->panner);
    std::vector<double> pv(pan_rows.size() * (size_t)S);
    rc = earb200_panner_handle_host(calc->panner, (int64_t)pan_rows.size(), pan_xyz.data(), pv.data(), nullptr,
                                        nullptr);
    if (rc != EARB200_OK) throw Error((Status)rc, last_error());
    calc->launches++;
    for (size_t k = 0; k < pan_rows.size(); k++) {
      double *g = gains + pan_rows[k] * C;
      int s = 0;
      for (int c = 0; c < C; c++)
        if (!calc->layout.channels[c].lfe) g[c] = pv[k * S + s++];
    }
  });
}

}  // extern "C"

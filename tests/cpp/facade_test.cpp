// Drives the C++14 facade (libear_b200/include/ear/*.hpp) the way an application drives libear.
// The scenarios follow libear's own tests (tests/gain_calculator_objects_tests.cpp:74-169,
// gain_calculator_direct_speakers_tests.cpp, gain_calculator_hoa_tests.cpp, gain_interpolator_tests.cpp,
// block_convolver_tests.cpp, delay_buffer_tests.cpp, variable_block_size_tests.cpp); written
// against the facade only.  `selftest` checks behaviour in-process; `objects <layout>` reads
// metadata rows from stdin and prints the gains so that pytest can compare them with the golden
// vectors of the compiled reference.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <numeric>
#include <sstream>

#include "ear/conversion.hpp"
#include "ear/dsp/dsp.hpp"
#include "ear/ear.hpp"

using namespace ear;

static int failures = 0;
#define CHECK(cond)                                                        \
  do {                                                                     \
    if (!(cond)) {                                                         \
      std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond);          \
      failures++;                                                          \
    }                                                                      \
  } while (0)
#define CHECK_THROWS(expr, type)                                           \
  do {                                                                     \
    bool caught = false;                                                   \
    try {                                                                  \
      expr;                                                                \
    } catch (const type &) {                                               \
      caught = true;                                                       \
    } catch (...) {                                                        \
    }                                                                      \
    if (!caught) {                                                         \
      std::printf("FAIL %s:%d: %s did not throw %s\n", __FILE__, __LINE__, #expr, #type); \
      failures++;                                                          \
    }                                                                      \
  } while (0)

static bool close(double a, double b, double tol = 1e-6) { return std::abs(a - b) <= tol; }

static void test_layouts() {
  CHECK(loadLayouts().size() == 10);
  Layout l = getLayout("4+5+0");
  CHECK(l.channels().size() == 10);
  CHECK(l.channelNames()[0] == "M+030");
  CHECK(l.withoutLfe().channels().size() == 9);
  CHECK((bool)l.indexForName("LFE1") && l.indexForName("LFE1").get() == 3);
  CHECK(!l.indexForName("nope"));
  CHECK(l.channelWithName("M+110").azimuthRange().first == 100.0);
  CHECK_THROWS(getLayout("7+4+0"), unknown_layout);
}

static void test_objects() {
  Layout layout = getLayout("0+5+0");
  GainCalculatorObjects gc(layout);
  ObjectsTypeMetadata otm;
  otm.position = PolarPosition(15.0, 0.0, 1.0);
  std::vector<double> direct(6), diffuse(6);
  gc.calculate(otm, direct, diffuse);
  // half way between M+030 and M+000 (tests/point_source_panner_tests.cpp:421-519)
  CHECK(close(direct[0], std::sqrt(0.5), 1e-9) && close(direct[2], std::sqrt(0.5), 1e-9));
  CHECK(direct[1] == 0.0 && direct[3] == 0.0 && direct[4] == 0.0 && direct[5] == 0.0);
  for (double d : diffuse) CHECK(d == 0.0);

  otm.diffuse = 0.25;
  otm.gain = 0.5;
  std::vector<float> df(6), ff(6);
  gc.calculate(otm, df, ff);
  CHECK(close(df[0], 0.5 * std::sqrt(0.5) * std::sqrt(0.75)) && close(ff[0], 0.5 * std::sqrt(0.5) * 0.5));

  // wrong output size (include/ear/helpers/output_gains.hpp:40-44)
  std::vector<float> small(5);
  CHECK_THROWS(gc.calculate(otm, small, ff), invalid_argument);

  // features libear refuses (tests/gain_calculator_objects_tests.cpp:134-169)
  ObjectsTypeMetadata bad;
  CHECK_THROWS(gc.calculate(bad, df, ff), not_implemented);  // default position is Cartesian
  bad.position = PolarPosition(0.0, 0.0, 1.0);
  gc.calculate(bad, df, ff);
  bad.cartesian = true;
  CHECK_THROWS(gc.calculate(bad, df, ff), not_implemented);
  bad.cartesian = false;
  bad.channelLock.flag = true;
  CHECK_THROWS(gc.calculate(bad, df, ff), not_implemented);
  bad.channelLock.flag = false;
  bad.objectDivergence = PolarObjectDivergence(0.5);
  CHECK_THROWS(gc.calculate(bad, df, ff), not_implemented);
  bad.objectDivergence = PolarObjectDivergence(0.0);
  bad.zoneExclusion.zones.push_back(PolarExclusionZone{0, 1, 0, 1, 0, 1, "z"});
  CHECK_THROWS(gc.calculate(bad, df, ff), not_implemented);
  bad.zoneExclusion.zones.clear();
  bad.screenRef = true;
  try {
    gc.calculate(bad, df, ff);
    CHECK(false);
  } catch (const not_implemented &e) {
    CHECK(std::string(e.what()) == "not implemented: screenRef");
  }

  // batch == single
  std::vector<ObjectsTypeMetadata> batch(64);
  for (size_t i = 0; i < batch.size(); i++) {
    batch[i].position = PolarPosition(-180.0 + 5.5 * i, -40.0 + 1.3 * i, 1.0);
    batch[i].width = 3.0 * i;
    batch[i].height = 1.0 * i;
    batch[i].diffuse = i / 64.0;
  }
  std::vector<float> bd, bf;
  gc.calculate_batch(batch, bd, bf);
  for (size_t i = 0; i < batch.size(); i += 9) {
    gc.calculate(batch[i], df, ff);
    for (int c = 0; c < 6; c++) CHECK(bd[i * 6 + c] == df[c] && bf[i * 6 + c] == ff[c]);
  }
}

static void test_direct_speakers() {
  Layout layout = getLayout("4+7+0");
  GainCalculatorDirectSpeakers gc(layout);
  std::vector<double> g(layout.channels().size());
  int warnings = 0;
  WarningCB count = [&](const Warning &) { warnings++; };

  DirectSpeakersTypeMetadata md;
  md.speakerLabels = {"urn:itu:bs:2051:0:speaker:M+030"};
  md.position = PolarSpeakerPosition(30.0, 0.0, 1.0);
  gc.calculate(md, g, count);
  CHECK(g[0] == 1.0 && std::accumulate(g.begin(), g.end(), 0.0) == 1.0 && warnings == 0);

  md.speakerLabels = {"LFE1"};  // LFE by name without a frequency element: mismatch warning
  gc.calculate(md, g, count);
  CHECK(g[layout.indexForName("LFE1").get()] == 1.0 && warnings == 1);
  md.channelFrequency.lowPass = 120.0;  // name and frequency agree
  warnings = 0;
  gc.calculate(md, g, count);
  CHECK(g[layout.indexForName("LFE1").get()] == 1.0 && warnings == 0);
  md.channelFrequency.lowPass = 500.0;  // present, but not an LFE frequency: both warnings
  std::vector<Warning::Code> codes;
  gc.calculate(md, g, [&](const Warning &w) { codes.push_back(w.code); });
  CHECK(codes.size() == 2 && codes[0] == Warning::Code::FREQ_NOT_LFE &&
        codes[1] == Warning::Code::FREQ_SPEAKERLABEL_LFE_MISMATCH);

  DirectSpeakersTypeMetadata pan;
  pan.speakerLabels = {"foo"};
  pan.position = PolarSpeakerPosition(12.0, 5.0, 1.0);
  gc.calculate(pan, g, count);
  double power = 0.0;
  for (double v : g) power += v * v;
  CHECK(close(power, 1.0, 1e-12) && g[layout.indexForName("LFE1").get()] == 0.0);

  PolarSpeakerPosition bounded(70.0, 28.0, 1.0);
  bounded.azimuthMin = 40.0;
  bounded.azimuthMax = 100.0;
  bounded.elevationMin = 20.0;
  bounded.elevationMax = 40.0;
  pan.position = bounded;
  gc.calculate(pan, g, count);
  CHECK(g[layout.indexForName("U+045").get()] == 1.0);

  DirectSpeakersTypeMetadata cart;
  cart.position = CartesianSpeakerPosition(0.0, 1.0, 0.0);
  CHECK_THROWS(gc.calculate(cart, g, count), not_implemented);
  DirectSpeakersTypeMetadata pack;
  pack.audioPackFormatID = std::string("AP_00010003");
  CHECK_THROWS(gc.calculate(pack, g, count), adm_error);
  std::vector<double> small(3);
  CHECK_THROWS(gc.calculate(md, small, count), invalid_argument);

  // mapping rule: U+030 of a 4+5+0 bed on 0+5+0 goes to M+030
  Layout l050 = getLayout("0+5+0");
  GainCalculatorDirectSpeakers gc050(l050);
  DirectSpeakersTypeMetadata rule;
  rule.speakerLabels = {"U+030"};
  rule.position = PolarSpeakerPosition(30.0, 30.0, 1.0);
  rule.audioPackFormatID = std::string("AP_00010005");
  std::vector<double> g5(6);
  gc050.calculate(rule, g5, count);
  CHECK(g5[0] == 1.0);

  GainCalculatorDirectSpeakers subs(layout, {{"left", "M+030"}});
  DirectSpeakersTypeMetadata s;
  s.speakerLabels = {"left"};
  subs.calculate(s, g, count);
  CHECK(g[0] == 1.0);
}

static void test_hoa() {
  Layout layout = getLayout("4+7+0");
  GainCalculatorHOA gc(layout);
  HOATypeMetadata md;
  for (int n = 0; n <= 3; n++)
    for (int m = -n; m <= n; m++) {
      md.orders.push_back(n);
      md.degrees.push_back(m);
    }
  std::vector<std::vector<double>> gains(16, std::vector<double>(layout.channels().size()));
  gc.calculate(md, gains);
  const size_t lfe = layout.indexForName("LFE1").get();
  for (auto &col : gains) CHECK(col[lfe] == 0.0);
  // the omnidirectional channel feeds every loudspeaker with the same sign
  for (size_t o = 0; o < layout.channels().size(); o++)
    if (o != lfe) CHECK(gains[0][o] > 0.0);
  // first-order Y (left-right) channel: left speakers positive, right negative
  CHECK(gains[1][layout.indexForName("M+090").get()] > 0.0 && gains[1][layout.indexForName("M-090").get()] < 0.0);

  int warnings = 0;
  md.screenRef = true;
  md.nfcRefDist = 1.0;
  gc.calculate(md, gains, [&](const Warning &) { warnings++; });
  CHECK(warnings == 2);
  md.normalization = "foo";
  CHECK_THROWS(gc.calculate(md, gains), adm_error);
  md.normalization = "N3D";
  md.degrees[3] = 5;
  CHECK_THROWS(gc.calculate(md, gains), invalid_argument);
  md.degrees[3] = 1;
  std::vector<std::vector<double>> wrong(15, std::vector<double>(layout.channels().size()));
  CHECK_THROWS(gc.calculate(md, wrong), invalid_argument);
}

static void test_gain_interpolator() {
  using namespace ear::dsp;
  // LinearInterpVector: one ramp, constant before and after (gain_interpolator_tests.cpp:58-130)
  GainInterpolator<LinearInterpVector> gi;
  gi.interp_points.emplace_back(100, std::vector<float>{0.0f, 1.0f});
  gi.interp_points.emplace_back(200, std::vector<float>{1.0f, 0.0f});
  const size_t T = 300;
  std::vector<float> in(T, 1.0f), o0(T), o1(T);
  const float *ins[1] = {in.data()};
  float *outs[2] = {o0.data(), o1.data()};
  gi.process(0, T, ins, outs);
  for (size_t t = 0; t < T; t++) {
    const float p = t < 100 ? 0.0f : (t >= 200 ? 1.0f : (float)(t - 100) / 100.0f);
    CHECK(close(o0[t], p, 1e-6) && close(o1[t], 1.0f - p, 1e-6));
  }
  // same curve processed in ragged pieces gives the same samples
  std::vector<float> p0(T), p1(T);
  size_t sizes[] = {7, 93, 1, 150, 49};
  size_t pos = 0;
  for (size_t n : sizes) {
    const float *pi[1] = {in.data() + pos};
    float *po[2] = {p0.data() + pos, p1.data() + pos};
    gi.process((SampleIndex)pos, n, pi, po);
    pos += n;
  }
  CHECK(pos == T);
  for (size_t t = 0; t < T; t++) CHECK(p0[t] == o0[t] && p1[t] == o1[t]);

  // a step: two points at the same sample
  GainInterpolator<LinearInterpSingle> step;
  step.interp_points.emplace_back(50, 0.25f);
  step.interp_points.emplace_back(50, 0.75f);
  float *so[1] = {o0.data()};
  step.process(0, 100, ins, so);
  CHECK(o0[49] == 0.25f && o0[50] == 0.75f);

  // matrix: 3 in, 2 out, constant
  GainInterpolator<LinearInterpMatrix> mat;
  mat.interp_points.emplace_back(0, std::vector<std::vector<float>>{{1.0f, 0.0f}, {0.0f, 1.0f}, {0.5f, 0.5f}});
  std::vector<float> a(64, 1.0f), b(64, 2.0f), c(64, 4.0f);
  const float *mi[3] = {a.data(), b.data(), c.data()};
  float *mo[2] = {o0.data(), o1.data()};
  mat.process(1000, 64, mi, mo);
  for (int t = 0; t < 64; t++) CHECK(o0[t] == 3.0f && o1[t] == 4.0f);

  // unsorted points are reported when process() walks over them (gain_interpolator.hpp:123-124)
  GainInterpolator<LinearInterpSingle> unsorted;
  unsorted.interp_points.emplace_back(100, 1.0f);
  unsorted.interp_points.emplace_back(50, 0.0f);
  unsorted.interp_points.emplace_back(150, 0.0f);
  CHECK_THROWS(unsorted.process(0, 200, ins, so), invalid_argument);
}

static void test_block_convolver() {
  using namespace ear::dsp::block_convolver;
  const size_t B = 64;
  Context ctx(B, get_fft_kiss<float>());
  // a delta at tap 70 is a 70-sample delay spanning two partitions
  std::vector<float> taps(100, 0.0f);
  taps[70] = 2.0f;
  Filter filt(ctx, taps.size(), taps.data());
  CHECK(filt.num_blocks() == 2);
  BlockConvolver conv(ctx, filt);
  const size_t K = 6;
  std::vector<float> x(K * B), y(K * B);
  for (size_t i = 0; i < x.size(); i++) x[i] = std::sin(0.05f * (float)i) + 0.001f * (float)(i % 13);
  for (size_t k = 0; k < K; k++) conv.process(x.data() + k * B, y.data() + k * B);
  for (size_t i = 0; i < x.size(); i++) CHECK(close(y[i], i >= 70 ? 2.0f * x[i - 70] : 0.0f, 2e-6));
  // the same in one call
  BlockConvolver conv2(ctx, filt);
  std::vector<float> y2(K * B);
  conv2.process_blocks(K, x.data(), y2.data());
  for (size_t i = 0; i < x.size(); i++) CHECK(close(y2[i], y[i], 1e-6));
  // silence in, then the tail comes out
  BlockConvolver conv3(ctx, filt);
  std::vector<float> t0(B), t1(B), t2(B);
  conv3.process(x.data(), t0.data());
  conv3.process(nullptr, t1.data());
  conv3.process(nullptr, t2.data());
  for (size_t i = 0; i < B; i++) {
    const size_t g = B + i;
    CHECK(close(t1[i], g - 70 < B ? 2.0f * x[g - 70] : 0.0f, 2e-6));
  }
  CHECK(close(t2[B - 1], 0.0f, 1e-6));
  // filter checks (block_convolver_impl.cpp:83-96)
  std::vector<float> longtaps(3 * B + 1, 0.0f);
  Filter toolong(ctx, longtaps.size(), longtaps.data());
  CHECK_THROWS(conv.set_filter(toolong), ear::invalid_argument);
  Context other(128, get_fft_kiss<float>());
  Filter otherf(other, taps.size(), taps.data());
  CHECK_THROWS(conv.crossfade_filter(otherf), ear::invalid_argument);
  // fade down: after the crossfade block and the filter length the output is silent
  conv.fade_down();
  std::vector<float> ones(B, 1.0f), out(B);
  for (int k = 0; k < 4; k++) conv.process(ones.data(), out.data());
  for (size_t i = 0; i < B; i++) CHECK(close(out[i], 0.0f, 1e-6));
}

static void test_fft_plugin() {
  FFTImpl<float> &impl = get_fft_kiss<float>();
  const size_t n = 256;
  std::shared_ptr<FFTPlan<float>> plan = impl.plan(n);
  std::unique_ptr<FFTWorkBuf> work = plan->alloc_workbuf();
  std::vector<float> x(n), back(n);
  for (size_t i = 0; i < n; i++) x[i] = std::cos(2.0 * M_PI * 5.0 * (double)i / (double)n) + 0.25f;
  std::vector<std::complex<float>> X(n / 2 + 1);
  plan->transform_forward(x.data(), X.data(), *work);
  CHECK(close(X[0].real(), 0.25 * n, 1e-3) && close(X[5].real(), 0.5 * n, 1e-3) && close(X[5].imag(), 0.0, 1e-3));
  CHECK(close(std::abs(X[6]), 0.0, 1e-3));
  plan->transform_reverse(X.data(), back.data(), *work);
  for (size_t i = 0; i < n; i++) CHECK(close(back[i], (float)n * x[i], 2e-4 * n));
}

static void test_delay_and_adapter() {
  using namespace ear::dsp;
  const size_t C = 2, T = 1000, D = 255;
  std::vector<float> x(C * T), y(C * T);
  for (size_t i = 0; i < x.size(); i++) x[i] = (float)i;
  DelayBuffer delay(C, D);
  CHECK(delay.get_delay() == (int)D);
  PtrAdapterConst in(C);
  PtrAdapter out(C);
  for (size_t pos = 0; pos < T;) {
    const size_t n = std::min<size_t>(T - pos, 1 + pos % 97);
    in.set_strided(x.data(), T, pos);
    out.set_strided(y.data(), T, pos);
    delay.process(n, in.ptrs(), out.ptrs());
    pos += n;
  }
  for (size_t c = 0; c < C; c++)
    for (size_t t = 0; t < T; t++) CHECK(y[c * T + t] == (t < D ? 0.0f : x[c * T + t - D]));

  // VariableBlockSizeAdapter around a fixed-block function: output = input delayed by block_size
  const size_t B = 128;
  int calls = 0;
  VariableBlockSizeAdapter adapter(B, C, C, [&](const float *const *i, float *const *o) {
    calls++;
    for (size_t c = 0; c < C; c++)
      for (size_t t = 0; t < B; t++) o[c][t] = 2.0f * i[c][t];
  });
  CHECK(adapter.get_delay() == (int)B);
  std::vector<float> z(C * T);
  for (size_t pos = 0; pos < T;) {
    const size_t n = std::min<size_t>(T - pos, 1 + (pos * 7) % 301);
    in.set_strided(x.data(), T, pos);
    out.set_strided(z.data(), T, pos);
    adapter.process(n, in.ptrs(), out.ptrs());
    pos += n;
  }
  CHECK(calls == (int)(T / B));
  for (size_t c = 0; c < C; c++)
    for (size_t t = 0; t < T; t++) CHECK(z[c * T + t] == (t < B ? 0.0f : 2.0f * x[c * T + t - B]));
}

static void test_decorrelators() {
  Layout layout = getLayout("0+5+0");
  auto filters = designDecorrelators<float>(layout);
  CHECK(filters.size() == 6 && filters[0].size() == 512);
  CHECK(decorrelatorCompensationDelay() == 255);
  auto one = designDecorrelator<double>(layout, 2);
  for (size_t i = 0; i < 512; i += 37) CHECK(close(one[i], filters[2][i], 1e-7));
  // all-pass: unit energy
  double e = 0.0;
  for (double v : one) e += v * v;
  CHECK(close(e, 1.0, 1e-3));
}

// ref tests/conversion_tests.cpp:52-90, 162-252
static void test_conversion() {
  using namespace ear::conversion;
  for (double az : {0.0, -10.0, 10.0, 90.0, -90.0, 150.0, -150.0})
    for (double el : {0.0, -10.0, 10.0, -45.0, 45.0, -90.0, 90.0})
      for (double dist : {0.5, 1.0}) {
        PolarPosition back = pointCartToPolar(pointPolarToCart(PolarPosition(az, el, dist)));
        if (std::abs(el) < 90.0) CHECK(close(back.azimuth, az));
        CHECK(close(back.elevation, el) && close(back.distance, dist));
      }
  {
    auto res = extentPolarToCart(PolarPosition(10.0, 20.0, 0.3), ExtentParams{40.0, 50.0, 0.6});
    CHECK(close(res.first.X, -0.08972503721988338) && close(res.first.Y, 0.3) &&
          close(res.first.Z, 0.1732050807568877));
    CHECK(close(res.second.width, 0.35166171614357594) && close(res.second.height, 0.4470181645863707) &&
          close(res.second.depth, 0.5762749096794243));
  }
  {
    auto res = extentCartToPolar(CartesianPosition(0.9, 0.8, 0.1), ExtentParams{0.3, 0.5, 0.4});
    CHECK(close(res.first.azimuth, -34.85107611658391) && close(res.first.elevation, 4.226794497273273) &&
          close(res.first.distance, 0.9000000000000001));
    CHECK(close(res.second.width, 76.50724453298275) && close(res.second.height, 104.9708107421662) &&
          close(res.second.depth, 0.1756348204517474));
  }
  // wrappers: the type held by position decides, the flag is fixed up
  ObjectsTypeMetadata otm;
  otm.position = PolarPosition(10.0, 20.0, 0.3);
  otm.width = 40.0, otm.height = 50.0, otm.depth = 0.6;
  ObjectsTypeMetadata otm_cart = otm;
  toCartesian(otm_cart);
  CHECK(otm_cart.cartesian && otm_cart.position.which() == 0);
  CHECK(close(compat::get<CartesianPosition>(otm_cart.position).Y, 0.3) && close(otm_cart.depth, 0.5762749096794243));
  otm.cartesian = true;
  toPolar(otm);
  CHECK(!otm.cartesian && close(otm.width, 40.0));
  toPolar(otm_cart);
  CHECK(!otm_cart.cartesian && otm_cart.position.which() == 1);
  // a Cartesian-authored block through the batched pre-pass, then the calculator
  std::vector<ObjectsTypeMetadata> block(3);
  block[0].position = CartesianPosition(0.9, 0.8, 0.1);
  block[0].cartesian = true;
  block[0].width = 0.3, block[0].height = 0.5, block[0].depth = 0.4;
  block[1].position = PolarPosition(30.0, 0.0, 1.0);
  block[2].position = CartesianPosition(-1.0, 1.0, 0.0);
  block[2].cartesian = true;
  GainCalculatorObjects gc(getLayout("0+5+0"));
  std::vector<double> direct(6), diffuse(6);
  CHECK_THROWS(gc.calculate(block[0], direct, diffuse), not_implemented);
  toPolar(block);
  CHECK(close(block[0].width, 76.50724453298275) && close(compat::get<PolarPosition>(block[2].position).azimuth, 30.0));
  gc.calculate(block[2], direct, diffuse);
  CHECK(close(direct[0], 1.0));  // M+030
  CHECK_THROWS(pointPolarToCart(PolarPosition(std::nan(""), 0.0, 1.0)), internal_error);
}

static int run_objects(const std::string &layout_name) {
  Layout layout = getLayout(layout_name);
  GainCalculatorObjects gc(layout);
  std::vector<ObjectsTypeMetadata> batch;
  std::string line;
  while (std::getline(std::cin, line)) {
    std::istringstream ss(line);
    double az, el, dist;
    ObjectsTypeMetadata m;
    if (!(ss >> az >> el >> dist >> m.width >> m.height >> m.depth >> m.gain >> m.diffuse)) continue;
    m.position = PolarPosition(az, el, dist);
    batch.push_back(m);
  }
  const size_t n_out = layout.channels().size();
  std::vector<double> direct(n_out), diffuse(n_out);
  for (const ObjectsTypeMetadata &m : batch) {
    gc.calculate(m, direct, diffuse);
    for (double v : direct) std::printf("%.17g ", v);
    for (double v : diffuse) std::printf("%.17g ", v);
    std::printf("\n");
  }
  return 0;
}

// OPT-IN extensions through libear's own metadata types: the default is libear's not_implemented;
// after enableExtensions a fully diverged point source straight ahead sits half and half (in
// power) on M+030 / M-030, and a locked object renders as the loudspeaker it snaps to
static void test_extensions() {
  Layout layout = getLayout("0+5+0");
  GainCalculatorObjects gc(layout);
  std::vector<double> direct(6), diffuse(6);
  ObjectsTypeMetadata m;
  m.position = PolarPosition(0.0, 0.0, 1.0);
  m.objectDivergence = PolarObjectDivergence(1.0, 30.0);
  bool threw = false;
  try {
    gc.calculate(m, direct, diffuse);
  } catch (const not_implemented &) {
    threw = true;
  }
  CHECK(threw);
  gc.enableExtensions(true, true);
  gc.calculate(m, direct, diffuse);
  const std::vector<std::string> names = layout.channelNames();
  for (size_t c = 0; c < names.size(); c++) {
    const double want = (names[c] == "M+030" || names[c] == "M-030") ? std::sqrt(0.5) : 0.0;
    CHECK(close(direct[c], want, 1e-6));
  }
  ObjectsTypeMetadata l;
  l.position = PolarPosition(-107.0, 2.0, 1.0);
  l.channelLock = ChannelLock(true, 0.2);
  gc.calculate(l, direct, diffuse);
  for (size_t c = 0; c < names.size(); c++) CHECK(close(direct[c], names[c] == "M-110" ? 1.0 : 0.0, 1e-9));
  // zoneExclusion stays refused
  ObjectsTypeMetadata z;
  z.position = PolarPosition(0.0, 0.0, 1.0);
  z.zoneExclusion.zones.push_back(PolarExclusionZone{0, 10, 0, 10, 0, 1, "zone"});
  threw = false;
  try {
    gc.calculate(z, direct, diffuse);
  } catch (const not_implemented &) {
    threw = true;
  }
  CHECK(threw);
}

// mode "vbs <block_size> <n_in> <n_out> <size>...": the facade's VariableBlockSizeAdapter over calls
// of the given sizes, around the same probe process function as oracle/ref_capi.cpp::ref_vbs_process
// runs libear's adapter with (ref src/dsp/variable_block_size_impl.cpp:44-81,
// tests/variable_block_size_tests.cpp).  stdin: float32 [n_in][T] raw, stdout: float32 [n_out][T] raw.
// Host-only: needs no device.
static int run_vbs(int argc, char **argv) {
  using namespace ear::dsp;
  const size_t B = std::stoul(argv[2]), n_in = std::stoul(argv[3]), n_out = std::stoul(argv[4]);
  std::vector<size_t> sizes;
  size_t T = 0;
  for (int a = 5; a < argc; a++) {
    sizes.push_back(std::stoul(argv[a]));
    T += sizes.back();
  }
  std::vector<float> x(n_in * T), y(n_out * T, -7.0f);
  if (std::fread(x.data(), sizeof(float), x.size(), stdin) != x.size()) return 3;
  int inner_calls = 0;
  VariableBlockSizeAdapter adapter(B, n_in, n_out, [&](const float *const *i, float *const *o) {
    for (size_t oc = 0; oc < n_out; oc++)
      for (size_t t = 0; t < B; t++) o[oc][t] = (float)(oc + 2) * i[oc % n_in][t] + 0.25f * (float)inner_calls;
    inner_calls++;
  });
  if (adapter.get_delay() != (int)B) return 4;
  PtrAdapterConst in(n_in);
  PtrAdapter out(n_out);
  size_t pos = 0;
  for (size_t n : sizes) {
    in.set_strided(x.data(), T, pos);
    out.set_strided(y.data(), T, pos);
    adapter.process(n, in.ptrs(), out.ptrs());
    pos += n;
  }
  std::fwrite(y.data(), sizeof(float), y.size(), stdout);
  return 0;
}

int main(int argc, char **argv) {
  try {
    if (argc >= 5 && std::string(argv[1]) == "vbs") return run_vbs(argc, argv);
    if (argc >= 3 && std::string(argv[1]) == "objects") return run_objects(argv[2]);
    test_layouts();
    test_decorrelators();
    test_objects();
    test_direct_speakers();
    test_hoa();
    test_gain_interpolator();
    test_block_convolver();
    test_fft_plugin();
    test_delay_and_adapter();
    test_conversion();
    test_extensions();
  } catch (const std::exception &e) {
    std::printf("FAIL unexpected exception: %s\n", e.what());
    return 2;
  }
  std::printf(failures ? "facade selftest: %d failure(s)\n" : "facade selftest: ok\n", failures);
  return failures ? 1 : 0;
}

// Launchers of the CUDA kernels (implemented in the *_kernels.cu files).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <vector>

#include "tables.h"


namespace earb {

  // Kernel attributes (the opt-in to more than 48 KB of dynamic shared memory) are per DEVICE:
  // a process may hold handles on several GPUs, so "configured once" is tracked per device.
  // Returns true the first time it is called for the current device with this mask.
  inline bool first_use_on_current_device(std::atomic<unsigned long long> &mask) {
    int dev = 0;
    cudaGetDevice(&dev);
    const unsigned long long bit = 1ull << (dev & 63);
    return !(mask.fetch_or(bit) & bit);
  }

  // device-resident struct-of-arrays form of ear::ObjectsTypeMetadata
  // (ref include/ear/metadata.hpp:127-145): the eight fields the polar path reads
  struct ObjectsSoA {
    const double *azimuth, *elevation, *distance;
    const double *width, *height, *depth;
    const double *gain, *diffuse;
    // optional [n][3] Cartesian positions used instead of azimuth/elevation/distance
    // (PolarExtent::handle takes the Cartesian position; ref polar_extent.hpp:62)
    const double *xyz;
  };

  // ---- opt-in extensions: divergence and channel lock (objects_ext_kernels.cu)
  void launch_objects_ext_expand(long n, const ObjectsSoA &in, const double *divergence, const double *az_range,
                                 const double *lock_max_distance, int n_spk, const double *d_spk, double *xyz3,
                                 double *w3, double *h3, double *d3, double *ones, double *zeros, double *weights,
                                 cudaStream_t stream);
  template <typename OutT>
  void launch_objects_ext_combine(long n, int n_out, const double *g3, const double *weights, const double *gain,
                                  const double *diffuse, OutT *out_direct, OutT *out_diffuse, cudaStream_t stream);

  // ---- K1 / K2: gain calculation (objects_kernels.cu)
  void launch_psp_batch(const PspTables *d_tables, long n, const double *d_xyz,
                        double *d_gains, int32_t *d_region, int32_t *d_inner,
                        cudaStream_t stream);
  // writes the active passes compacted to d_passes[0 .. d_counters[0]) and the objects that need the
  // point source panner (all of them with all_points) to d_point_list[0 .. d_counters[1]); both
  // counters must be zero (thresholds: also the fade's start / end cos and sin, which only the FFMA
  // core reads)
  void launch_objects_prelude(long n, const ObjectsSoA &in, ExtentPass *d_passes, ObjectAux *d_aux,
                              unsigned int *d_counters, int32_t *d_point_list, bool all_points,
                              cudaStream_t stream);
  // scratch of earb200_objects_calculate_device_static (objects_static.cu): a file-scope device array,
  // no cudaMalloc.  Layout: [counters, 256 B][passes 2N][aux N][results 2N x 24][point gains N x 24]
  // [point list N], N = EARB_OBJECTS_STATIC_MAX objects per call.
#define EARB_OBJECTS_STATIC_MAX 4096
#define EARB_OBJECTS_STATIC_BYTES \
  (256 + EARB_OBJECTS_STATIC_MAX * (2 * 96 + 24 + 2 * EARB_MAX_SPK * 4 + EARB_MAX_SPK * 8 + 4))
  static_assert(sizeof(ExtentPass) == 96 && sizeof(ObjectAux) == 24, "static scratch layout");
  // device address of that array on `device` (which must be current); looked up once per device
  cudaError_t objects_static_scratch(int device, void **ptr, size_t *bytes);
#define EXTENT_TC_STAGE_POINTS 64
  // tcgen05 extent core (extent_tc_kernels.cu): thread-per-pass weights feeding 3xTF32 UMMAs
  size_t extent_tc_smem_bytes(const ExtentGrid &grid);
  size_t extent_tc_bimage_floats(const ExtentGrid &grid);
  void extent_tc_build_bimage(const ExtentGrid &grid, const float *gains, float *image);
  void launch_extent_core_tc(const ExtentGrid &grid, const float4 *d_points, const float *d_bimage,
                             const unsigned int *d_n_active, const ExtentPass *d_passes,
                             float *d_results, int n_sms, cudaStream_t stream);
  // cell-culled tcgen05 extent core (extent_cell_kernels.cu): the virtual sources clustered into
  // compact cells of 16, uniform (pass, cell) pairs decided from the cell centre, the rest
  // evaluated in batches; fp16 hi/lo UMMAs with the weights staged in shared memory
#define EXTENT_CELL_POINTS 16
  struct ExtentCells {
    int n_cells = 0;
    std::vector<int> order;      // [n_cells][16] grid point of every slot, -1: padding
    std::vector<float> centre;   // [n_cells][3] bounding-cap centre as the kernel sees it
    std::vector<double> radius;  // [n_cells] bounding-cap radius, radians
  };
  // xyz: [3][n_points_pad] float (the grid as the reference's float core sees it)
  void extent_cells_build(int n_points, const float *xyz, int n_points_pad, ExtentCells &out);
  size_t extent_cell_stream_bytes(const ExtentCells &cells);
  size_t extent_cell_centre_floats(const ExtentCells &cells);
  size_t extent_cell_sum_bytes(const ExtentCells &cells);
  size_t extent_cell_point_floats(const ExtentCells &cells);
  // gains: speaker-major [S][Pp] float in grid order; xyz: [3][Pp] float
  void extent_cell_build_stream(const ExtentCells &cells, int n_spk, int n_points_pad, const float *gains,
                                unsigned char *stream);
  void extent_cell_build_sums(const ExtentCells &cells, int n_spk, int n_points_pad, const float *gains,
                              unsigned char *sums);
  void extent_cell_build_points(const ExtentCells &cells, int n_points_pad, const float *xyz, float *out);
  void extent_cell_build_centres(const ExtentCells &cells, float *out);
  size_t extent_cell_smem_bytes(const ExtentCells &cells);
  bool extent_cell_supported(const ExtentCells &cells);
  void launch_extent_cell(int n_cells, int n_spk, const unsigned char *d_stream, const unsigned char *d_sums,
                          const float *d_centres, const float *d_points, const unsigned int *d_n_active,
                          const ExtentPass *d_passes, float *d_results, int n_sms, cudaStream_t stream);
  // self-test: counts the elements where the epilogue's branch-free square root differs from sqrt()
  void launch_selftest_fast_sqrt(const double *d_x, long n, unsigned long long *d_n_mismatch, cudaStream_t stream);
  // two launches: the point source gains of the objects that need them (d_point_gains
  // [n][EARB_MAX_SPK] scratch, and the region indices when d_region is given), then the gains
  template <typename OutT>
  void launch_objects_epilogue(const PspTables *d_tables, int n_real, long n, const ObjectsSoA &in,
                               const ExtentPass *d_passes, const ObjectAux *d_aux,
                               const float *d_results, double *d_point_gains, const unsigned int *d_counters,
                               const int32_t *d_point_list, int n_sms, int n_out,
                               const int32_t *d_out_index, OutT *d_direct, OutT *d_diffuse,
                               int32_t *d_region, cudaStream_t stream);

  // ---- K3: interpolated mixing (render_kernels.cu)
#define RENDER_THREADS 128
#define RENDER_UNROLL 4
  // one interpolation segment of the call: gains ramp from point pa to point pb
  // between ramp_start and ramp_start + 1/scale (sample offsets relative to the start of
  // the call, may be negative), or stay at point pa (ramp == 0)
  struct RenderSegment {
    long long ramp_start;
    float scale;  // 1.0f / (end - start)   (ref gain_interpolator.hpp:255)
    int pa, pb;
    int ramp;
  };
  // a run of samples [t0, t0 + len) of the call inside one segment
  struct RenderItem {
    long long t0;
    int len;
    int seg;
  };
#define RENDER_PIPE_TILE 512
#define RENDER_TC_TILE 256
  // work items of one call, grouped: full 16-byte aligned RENDER_PIPE_TILE-sample tiles go to
  // the TMA-staged kernel ("fast"), ragged / unaligned runs to the direct kernel ("slow")
  // Loudspeaker bus shared by the GPUs of one node (csrc/capi_bus.cpp): the bus is cut into one
  // time slice per rank; a rank's render stores every output tile straight into the slot it has
  // in the staging memory of the slice's OWNER -- peer stores over NVLink from the kernel's
  // epilogue -- and the owner sums the slots.  n_ranks == 0: no bus, plain `out`.
#define EARB_BUS_MAX_RANKS 8
  struct RenderBus {
    float *slot[EARB_BUS_MAX_RANKS];  // per owner: this rank's slot there, [M][slice_len] floats
    long long slice_len;              // samples per slice
    long long t_origin;               // bus sample of the call's sample 0
    int n_ranks;
    int rank;  // rank r starts its walk over the tiles r / n_ranks of the way in: at any moment the
               // ranks store into DIFFERENT owners (all ranks walking the time axis in step would
               // all store into the one GPU that owns the current slice: 8-to-1 incast on its links)
  };
  // A second interpolator riding the same launch of the tcgen05 kernel: same input, same tiles and
  // segments, its own gain images and output rows (the direct and the diffuse render of the Objects
  // signal flow).  Its work units follow the first one's of the same tile, on the neighbouring CTA
  // at the same time, so the tile's input comes from HBM once and from L2 the second time.
  struct RenderPair {
    const float *images2 = nullptr;
    float *out2 = nullptr;  // same row pitch as `out`
  };
  struct RenderPlan {
    const RenderItem *d_items;  // [fast ramp | fast const | slow ramp | slow const]
    long n_fast_ramp, n_fast_const, n_slow_ramp, n_slow_const;
    int use_tc;  // fast items are RENDER_TC_TILE-sample tiles for the tcgen05 kernel
    int n_sms;   // grid size of the persistent tcgen05 kernel
    long n_samples;  // samples per channel of this call (extent of the TMA descriptor)
    const float *d_images;  // per-point gain-matrix images (tcgen05 path)
    int accumulate;         // out += result instead of out = result (mixing channel groups)
    long n_points;          // interpolation points behind d_images
    RenderBus bus;          // n_ranks > 0: results go to the bus slots (accumulate still READS `out`)
  };
  // tcgen05 / TMEM variant of the fast path (render_tc_kernels.cu)
  bool render_tc_supported(int N, int M);
  size_t render_tc_smem_bytes(int N, int M);
  // d_items: n_tiles tiles of RENDER_TC_TILE samples, ramp and constant segments alike
  // (the input is read through a 2-D TMA descriptor over [N][n_samples]); false = descriptor
  // could not be encoded, nothing launched
  bool launch_render_tc(int N, int M, const float *in, long in_stride, long n_samples, float *out,
                        long out_stride, const RenderItem *d_items, long n_tiles,
                        const RenderSegment *d_segs, const float *d_images, long n_points, int n_sms,
                        int accumulate, const RenderBus &bus, cudaStream_t stream,
                        const RenderPair &pair = RenderPair());
  // per-point gain-matrix images for the tcgen05 kernel (TF32 hi/lo split in UMMA layout)
  size_t render_tc_image_floats(int N, int M);
  void launch_render_prepare_points(int N, int M, long n_points, const float *d_gains, float *d_images,
                                    cudaStream_t stream);
  // d_flags[i] (i >= 1) = gain matrix i equals matrix i - 1 (mat floats each)
  void launch_render_points_same(long mat, long n_points, const float *d_gains, char *d_flags, cudaStream_t stream);
  int render_tile_samples(int M);
  size_t render_smem_bytes(int N, int M);
  void launch_render(int N, int M, const float *in, long in_stride, float *out, long out_stride,
                     const RenderPlan &plan, const RenderSegment *d_segs, const float *d_gains,
                     cudaStream_t stream);

  // ---- bus exchange between the GPUs of a node (bus_kernels.cu)
  // flags[src] of every peer <- epoch (release at system scope): this rank's stores of the epoch
  // into the peers' slots are complete
  void launch_bus_signal(uint32_t *const *peer_flags, int n_ranks, int rank, uint32_t epoch, cudaStream_t stream);
  // waits until flags[r] >= epoch for every rank, then out[m][t] = sum over r of slot r (own slice);
  // gather != null: also stores the summed slice into every rank's full bus
  void launch_bus_reduce(const uint32_t *flags, int n_ranks, uint32_t epoch, const float *slots, long long slot_floats,
                         int M, long long slice_len, long long valid_len, float *out, long long out_stride,
                         float *const *gather, long long gather_stride, long long gather_offset, int n_sms,
                         cudaStream_t stream);
  // local [M][T] partial bus -> the slots (the path for calls the fused epilogue cannot take)
  void launch_bus_push(const float *partial, long long partial_stride, int M, long long t_origin, long long n,
                       const RenderBus &bus, cudaStream_t stream);
  // the same for a list of runs (the items of a call the fused epilogue did not take)
  void launch_bus_push_items(const RenderItem *d_items, long n_items, const float *partial, long long partial_stride,
                             int M, const RenderBus &bus, cudaStream_t stream);
  // waits until flags[r] >= epoch for every rank (the gather of an epoch has landed everywhere)
  void launch_bus_wait(const uint32_t *flags, int n_ranks, uint32_t epoch, cudaStream_t stream);

  // mem_new[c][j] = sample n + j of [mem; in]: the delay line's memory after n samples, without
  // producing output (the delayed signal was consumed elsewhere: the convolver's fused store)
  void launch_delay_advance(int C, long n, long delay, const float *in, long in_stride, const float *mem,
                            float *mem_new, cudaStream_t stream);

  // ---- K4: partitioned FFT convolver bank (convolver_kernels.cu)
#define CONV_THREADS 128
  struct ConvLaunch {
    int B, nb, C, R, Kg;
    long K;
    const float *in;  // may be null (silence)
    long in_stride;
    float *out;
    long out_stride;
    const float *hist;      // [C][nb][B]
    float *hist_new;        // [C][nb][B]
    const float *tail;      // [C][B]
    float *tail_new;        // [C][B]
    const float *filt;      // [n_filters][nb][B+1] complex
    const short *ftab;      // [C][Kg][nb][2]
    const int *uniform;     // [C]
    const unsigned char *faded;  // [C][nb + Kg]
    const float *tw2;       // [2B] complex: exp(-2 pi i k / 2B)
    // fast path only: out = conv(in) + delay(add): the delayed direct bus of the Objects signal flow
    // joins the decorrelated diffuse bus in the convolver's store (ref docs/dsp.rst:40-71);
    // add [C][K * B] with row pitch add_stride, add_mem [C][add_delay] the delay line's memory
    const float *add = nullptr;
    long add_stride = 0, add_delay = 0;
    const float *add_mem = nullptr;
    // fast path with B * nb <= 512 other than one partition of 512: the whole filter as ONE 512-tap
    // partition -- filt_fast [n_filters][513] complex (1024-point spectrum of the nb * B taps), and
    // the taps themselves [n_filters][nb * B] for the carried tail (launch_conv_generic_tail)
    const float *filt_fast = nullptr;
    const float *taps = nullptr;
    // fast path, long calls (convolver_wide_kernel): filt_wide [n_filters][2048] complex, the
    // 2048-point spectrum of the whole filter; null: the 1024-point kernel takes the call
    const float *filt_wide = nullptr;
    int wide = 1;  // 0: never, 1: when the call is long enough to fill the GPU with long runs, 2: whenever possible
  };
  size_t conv_smem_bytes(int B, int nb);
  void launch_convolver(const ConvLaunch &L, cudaStream_t stream);
  void launch_conv_history(const ConvLaunch &L, cudaStream_t stream);  // hist -> hist_new only
  // fast path (convolver_fast_kernels.cu): B * nb <= 512 taps, no filter change in flight, non-null
  // input, K * B a multiple of 512; warp-per-pair-of-512-samples register FFTs.  d_filter_of: [C]
  // filter id per channel.  With B != 512 the carried state must be consistent with the current
  // filter (no set_filter since the last processed block) and launch_conv_generic_tail must follow.
  bool conv_fast_supported(int B, int nb);
  size_t conv_fast_tw_floats();
  void conv_fast_build_tw(float *tw);
  void launch_convolver_fast(const ConvLaunch &L, const int *d_filter_of, const float *d_tw_fast, int n_sms,
                             cudaStream_t stream);
  // tail_new [C][B] (unnormalised, as the generic kernel carries it) from hist_new and the taps of
  // each channel's filter: what libear's last_tail holds after the last block of the call
  void launch_conv_generic_tail(const ConvLaunch &L, const int *d_filter_of, cudaStream_t stream);

  // batched 2B-point real FFTs (the FFT plug-in seam): in [batch][2B] -> out [batch][B+1] complex
  // and back (unnormalised); tw2 = exp(-2 pi i k / 2B), k in [0, 2B)
  void launch_fft_r2c(int B, long batch, const float *in, float *out, const float *tw2, int n_sms,
                      cudaStream_t stream);
  void launch_fft_c2r(int B, long batch, const float *in, float *out, const float *tw2, int n_sms,
                      cudaStream_t stream);

  // ---- delay line / bus mix of the Objects signal flow (dsp_kernels.cu)
  // [mem; in] -> [out; mem_new] per channel (mem, mem_new: [C][delay], distinct buffers);
  // in may be null (silence)
  // add (nullable): out = delayed input + add (the last stage of the Objects signal flow)
  void launch_delay(int C, long n, long delay, const float *in, long in_stride, float *out,
                    long out_stride, const float *mem, float *mem_new, const float *add, long add_stride,
                    int n_sms, cudaStream_t stream);
  void launch_add(int C, long n, const float *a, long a_stride, const float *b, long b_stride, float *out,
                  long out_stride, int n_sms, cudaStream_t stream);

  // ---- Cartesian <-> polar metadata conversion pre-pass (conversion_kernels.cu)
  // the five azimuth sectors of ear::conversion (ref src/conversion.cpp:24-73): sector s runs
  // from corner s to corner (s + 1) % 5
  struct ConversionSectors {
    double polar_az[5];     // polar azimuth of the corners
    double cart_az[5];      // azimuth of the corners' Cartesian positions
    double cart_pos[5][2];  // Cartesian x, y of the corners
    double m[5][4];         // inverse of [start; end] per sector, row-major 2x2
  };
  // pos / ext: [n][3]; ext / ext_out nullable (positions only); *d_n_failed counts objects the
  // reference would have thrown on (their outputs are NaN)
  void launch_conversion(const ConversionSectors &sectors, long n, bool to_polar, const double *d_pos,
                         const double *d_ext, double *d_pos_out, double *d_ext_out,
                         unsigned int *d_n_failed, int n_sms, cudaStream_t stream);

  // ---- K5: HOA decoder design (hoa_kernels.cu)
  int hoa_power_blocks(int P);
  // Y [C][P], D [S][C], partial [hoa_power_blocks(P)]
  void launch_hoa_design(int P, int S, int C, const double *d_xyz, const double *d_G,
                         const int *d_orders, const int *d_degrees, const double *d_norms,
                         double *d_Y, double *d_D, double *d_partial, cudaStream_t stream);

}  // namespace earb

/* earb200.h -- C ABI of libear-b200: libear's data-parallel rendering hot path on
 * NVIDIA B200 (sm_100a).
 *
 * libear (ebu/libear) has no C ABI; its seams are C++ classes.  Each entry point
 * below names the libear interface it replaces (file:line under the libear tree).
 * The C++ facade in libear_b200/include/ear/ re-implements those classes with their
 * original signatures on top of these functions, and INTEGRATION.md shows the
 * binding a libear maintainer would add.
 *
 * Conventions
 *  - every function returns an earb200_status; on failure earb200_last_error()
 *    (thread-local) holds the message.  Codes 1..5 map 1:1 onto libear's exception
 *    classes (include/ear/exceptions.hpp:8-43), which the C++ facade re-throws.
 *  - handles are opaque, bound to one CUDA device, and -- like the libear objects
 *    they replace -- not thread-safe per instance.
 *  - *_host entry points take host buffers, copy, run and synchronise;
 *    *_device entry points take device pointers plus a cudaStream_t (passed as
 *    void*) and are asynchronous.  Buffers are caller-owned.
 *  - there is no CPU fallback: without a CUDA device every create call fails with
 *    EARB200_CUDA_ERROR.
 */
#ifndef EARB200_H
#define EARB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum earb200_status {
  EARB200_OK = 0,
  EARB200_NOT_IMPLEMENTED = 1,  /* ear::not_implemented */
  EARB200_INTERNAL_ERROR = 2,   /* ear::internal_error  */
  EARB200_ADM_ERROR = 3,        /* ear::adm_error       */
  EARB200_UNKNOWN_LAYOUT = 4,   /* ear::unknown_layout  */
  EARB200_INVALID_ARGUMENT = 5, /* ear::invalid_argument */
  EARB200_CUDA_ERROR = 6        /* CUDA runtime failure / no device */
} earb200_status;

const char *earb200_last_error(void);
const char *earb200_version(void);

/* ---------------------------------------------------------------- layouts
 * replaces ear::getLayout / ear::Layout (include/ear/bs2051.hpp:9,
 * include/ear/layout.hpp:15-96, src/bs2051.cpp:13-22). */
typedef struct earb200_layout earb200_layout;

int earb200_num_layouts(void);
const char *earb200_layout_name_at(int i);
/* a tabulated ITU-R BS.2051 layout by name, e.g. "9+10+3" */
int earb200_layout_create(const char *name, earb200_layout **out);
/* a BS.2051 layout with real loudspeaker positions that differ from the nominal
 * ones: `name` selects the tabulated layout (channel set, nominal positions);
 * real_pos is [n][3] = azimuth, elevation, distance per channel, table order. */
int earb200_layout_create_with_positions(const char *name, int n_channels,
                                         const double *real_pos, earb200_layout **out);
/* any layout (ear::Layout(name, channels), include/ear/layout.hpp:67-96): channels as given.
 * A name from the BS.2051 table selects that layout's tabulated hull facets; any other name
 * makes the panner triangulate the nominal positions itself (convex hull,
 * src/common/convex_hull.cpp:7-100, src/common/point_source_panner.cpp:484-487).
 * pos / nominal_pos [n][3] = azimuth, elevation, distance; az_range / el_range [n][2];
 * nominal_pos, az_range, el_range and is_lfe are nullable (defaults as ear::Channel:
 * nominal = real position, ranges = the position itself, not LFE). */
int earb200_layout_create_custom(const char *name, int n_channels,
                                 const char *const *channel_names, const double *pos,
                                 const double *nominal_pos, const double *az_range,
                                 const double *el_range, const int32_t *is_lfe,
                                 earb200_layout **out);
/* the triangulation used for such layouts, exposed for inspection (host, setup-time): facets of
 * the convex hull of n_points unit vectors xyz [n][3], in the order ear::convex_hull finds them;
 * facets is a flat list, each facet = vertex count followed by its sorted vertex indices. */
int earb200_convex_hull(int n_points, const double *xyz, int32_t *facets, int capacity,
                        int *n_written);
/* the clustering of the extent panner's virtual sources into cells of 16 that the extent core
 * culls by (K2; the reference's per-batch all-ones / all-zeros shortcut,
 * src/object_based/polar_extent_simd.hpp:111-126, decided per cell), exposed for inspection (host,
 * setup-time): grid = the reference's generatePanningPositionsEven(n_rows)
 * (polar_extent.cpp:15-39).  order [n_cells][16]: grid point of every slot (-1: padding),
 * centre [n_cells][3] and radius [n_cells] (radians): the bounding cap of each cell.  Buffers sized
 * for `capacity` cells; *n_cells_out / *n_points_out receive the counts. */
int earb200_extent_cells(int n_rows, int capacity, int32_t *order, float *centre, double *radius,
                         double *grid_xyz, int *n_cells_out, int *n_points_out);
void earb200_layout_destroy(earb200_layout *layout);
int earb200_layout_num_channels(const earb200_layout *layout);
const char *earb200_layout_channel_name(const earb200_layout *layout, int i);
int earb200_layout_channel_is_lfe(const earb200_layout *layout, int i);
/* pos[3] = azimuth, elevation, distance (real position); nominal likewise */
int earb200_layout_channel_position(const earb200_layout *layout, int i, double *pos,
                                    double *nominal);
/* allowed azimuth range (anticlockwise from [0] to [1]) and elevation range, degrees */
int earb200_layout_channel_ranges(const earb200_layout *layout, int i, double *az_range,
                                  double *el_range);

/* ------------------------------------------------- point-source panner (K1)
 * replaces ear::configurePolarPanner(layout.withoutLfe())->handle(position)
 * (src/common/point_source_panner.cpp:586-600, 234-244, 201-211).  Besides the S
 * gains it reports which region accepted the position: the index into the region
 * list configureFullPolarPanner builds (:495-554; virtual n-gons first, then the
 * hull facets in table order) and, for n-gons, the inner triplet index. */
typedef struct earb200_panner earb200_panner;

int earb200_panner_create(const earb200_layout *layout, int device, earb200_panner **out);
void earb200_panner_destroy(earb200_panner *panner);
int earb200_panner_num_channels(const earb200_panner *panner); /* S (no LFE) */
int earb200_panner_num_regions(const earb200_panner *panner);
/* xyz [n][3]; gains [n][S] (double); region_idx / inner_idx [n], nullable */
int earb200_panner_handle_host(earb200_panner *panner, int64_t n, const double *xyz,
                               double *gains, int32_t *region_idx, int32_t *inner_idx);
int earb200_panner_handle_device(earb200_panner *panner, int64_t n, const double *d_xyz,
                                 double *d_gains, int32_t *d_region_idx,
                                 int32_t *d_inner_idx, void *stream);

/* ------------------------------------------- Objects gain calculator (K1+K2)
 * replaces ear::GainCalculatorObjects (include/ear/gain_calculators.hpp:37-64,
 * src/object_based/gain_calculator_objects.cpp:26-57) for a batch of objects. */
typedef struct earb200_objects earb200_objects;

/* struct-of-arrays form of ear::ObjectsTypeMetadata (include/ear/metadata.hpp:127-145):
 * polar position (degrees / normalised distance), extent (degrees), gain, diffuse.
 * `flags` (nullable) carries the features libear refuses with ear::not_implemented
 * (gain_calculator_objects.cpp:37-43); a batch containing any of them fails the
 * same way. */
typedef struct earb200_objects_soa {
  const double *azimuth;
  const double *elevation;
  const double *distance;
  const double *width;
  const double *height;
  const double *depth;
  const double *gain;
  const double *diffuse;
  const int32_t *flags;
} earb200_objects_soa;

enum {
  EARB200_OBJ_CARTESIAN = 1,          /* metadata.cartesian == true           */
  EARB200_OBJ_CHANNEL_LOCK = 2,       /* channelLock.flag                     */
  EARB200_OBJ_SCREEN_REF = 4,         /* screenRef                            */
  EARB200_OBJ_DIVERGENCE = 8,         /* objectDivergence.divergence != 0     */
  EARB200_OBJ_ZONE_EXCLUSION = 16,    /* zoneExclusion.zones non-empty        */
  EARB200_OBJ_CARTESIAN_POSITION = 32 /* position holds a CartesianPosition   */
};

int earb200_objects_create(const earb200_layout *layout, int device, earb200_objects **out);
/* extension: the virtual-source grid of the extent panner.  EARB200_GRID_LIBEAR = libear's own
 * 1,652-point grid (generatePanningPositionsEven(37), src/object_based/polar_extent.cpp:12-39, 104;
 * what earb200_objects_create uses and all parity rests on); EARB200_GRID_TDESIGN_5200 = the
 * 5,200-point t-design of resources/Design_5200_100_random.dat: a throughput shape only -- libear
 * has no extent panner on that grid, so there is no reference output (parity unpinned). */
enum { EARB200_GRID_LIBEAR = 0, EARB200_GRID_TDESIGN_5200 = 1 };
int earb200_objects_create_with_grid(const earb200_layout *layout, int device, int grid,
                                     earb200_objects **out);
int earb200_objects_grid_points(const earb200_objects *calc);
void earb200_objects_destroy(earb200_objects *calc);
int earb200_objects_num_outputs(const earb200_objects *calc); /* C_out incl. LFE */
/* direct / diffuse: [n][C_out] row-major, float (out_is_double = 0) or double;
 * region_idx [n] nullable: region of the point-source panner at the object position.
 * In the _device variant the SoA pointers (including flags) are HOST pointers for
 * flags and DEVICE pointers for the eight double arrays. */
int earb200_objects_calculate_host(earb200_objects *calc, int64_t n,
                                   const earb200_objects_soa *in, int out_is_double,
                                   void *direct, void *diffuse, int32_t *region_idx);
int earb200_objects_calculate_device(earb200_objects *calc, int64_t n,
                                     const earb200_objects_soa *d_in, int out_is_double,
                                     void *d_direct, void *d_diffuse,
                                     int32_t *d_region_idx, void *stream);
/* earb200_objects_calculate_device for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments, same results, same kernels, but the
 * per-call scratch is a file-scope device array (csrc/objects_static.cu, 2,474,240 bytes, held by
 * every GPU that loads the library until the process exits) instead of buffers the handle grows on
 * demand, and all four kernels are queued on `stream` itself: the call makes no cudaMalloc / cudaFree
 * and uses no allocating library routine.  n > EARB200_OBJECTS_STATIC_MAX fails with
 * EARB200_INVALID_ARGUMENT.  Every call on a device shares the array (whatever the handle): calls
 * that overlap on the GPU -- two streams, two threads -- race; order them on one stream or with
 * events.  The first call on a device looks the array's address up (cudaGetSymbolAddress); make it
 * outside stream capture. */
#define EARB200_OBJECTS_STATIC_MAX 4096
int earb200_objects_calculate_device_static(earb200_objects *calc, int64_t n,
                                            const earb200_objects_soa *d_in, int out_is_double,
                                            void *d_direct, void *d_diffuse,
                                            int32_t *d_region_idx, void *stream);
/* OPT-IN extensions beyond libear (PARITY UNPINNED).  libear 0.9 rejects objectDivergence and
 * channelLock with ear::not_implemented (gain_calculator_objects.cpp:37-43; its tests require the
 * throws) and so does this library -- unless the caller switches them on here.  The arithmetic
 * then restates Rec. ITU-R BS.2127 as the EBU ADM Renderer implements it (csrc/objects_ext_kernels.cu);
 * there is no libear output to compare it with.  zoneExclusion, screenRef / screen edge lock and
 * Cartesian panning stay not_implemented. */
enum { EARB200_EXT_DIVERGENCE = 1, EARB200_EXT_CHANNEL_LOCK = 2 };
int earb200_objects_enable_extensions(earb200_objects *calc, uint32_t mask);
typedef struct earb200_objects_ext {
  const double *divergence;                /* objectDivergence.divergence in [0, 1]; NULL: none */
  const double *divergence_azimuth_range;  /* objectDivergence.azimuthRange; NULL: 45 degrees   */
  const double *channel_lock_max_distance; /* channelLock.maxDistance; < 0: flag off, +inf: no
                                              limit; NULL: no object is locked                  */
} earb200_objects_ext;
/* device arrays, asynchronous on `stream`; a feature that is present but not enabled fails with
 * EARB200_NOT_IMPLEMENTED and libear's message */
int earb200_objects_calculate_ext_device(earb200_objects *calc, int64_t n, const earb200_objects_soa *d_in,
                                         const earb200_objects_ext *d_ext, int out_is_double,
                                         void *d_direct, void *d_diffuse, void *stream);
/* the same with host arrays (copies, runs, synchronises) */
int earb200_objects_calculate_ext_host(earb200_objects *calc, int64_t n, const earb200_objects_soa *in,
                                       const earb200_objects_ext *ext, int out_is_double, void *direct,
                                       void *diffuse);
/* PolarExtent::handle on Cartesian positions (src/object_based/polar_extent.cpp:281-302):
 * xyz [n][3], width/height/depth [n]; out [n][S] double.  Host buffers. */
int earb200_objects_extent_host(earb200_objects *calc, int64_t n, const double *xyz,
                                const double *width, const double *height,
                                const double *depth, double *out);
/* number of kernels this handle has launched so far (bench bookkeeping) */
int64_t earb200_objects_launch_count(const earb200_objects *calc);
/* self-test hook (no reference counterpart): the gain epilogue takes its FP64 square roots on a
 * branch-free path that must equal sqrt() bit for bit; this runs that path over x[n] (host) on
 * `device` and returns the number of elements that differ (arguments outside the fast range fall
 * back to sqrt() in the product and are not counted). */
int earb200_selftest_fast_sqrt(int device, int64_t n, const double *x, int64_t *n_mismatch);
/* measurement hooks: with profiling on, every launch of the dominant kernel (the extent
 * core) is bracketed by CUDA events on the launching stream; _kernel_time synchronises,
 * returns the summed device time [ms] and launch count since the last query and the
 * number of pass slots (two per object; the second is idle when depth == 0) those
 * launches covered, then resets. */
int earb200_objects_set_profiling(earb200_objects *calc, int enable);
int earb200_objects_kernel_time(earb200_objects *calc, double *ms, int64_t *launches,
                                int64_t *passes);

/* -------------------------------------------------- gain interpolator (K3)
 * replaces ear::dsp::GainInterpolator<LinearInterpMatrix|Vector|Single>
 * (include/ear/dsp/gain_interpolator.hpp:40-130, 183-300): N input channels are
 * mixed into M outputs with piecewise-linear gains between interpolation points. */
typedef struct earb200_interp earb200_interp;

int earb200_interp_create(int n_in, int n_out, int device, earb200_interp **out);
void earb200_interp_destroy(earb200_interp *interp);
/* interp_points: sample index + [n_in][n_out] gain matrix each (host buffers).
 * Unsorted sample indices are reported lazily by process, as in libear (:123-124). */
int earb200_interp_set_points(earb200_interp *interp, int64_t n_points,
                              const int64_t *sample_idx, const float *gains);
/* the same with the gain matrices already in device memory -- [n_points][n_in][n_out] floats, e.g.
 * the `direct` rows earb200_objects_calculate_device wrote for n_in objects at each point: the
 * hand-off GainCalculatorObjects -> GainInterpolator::interp_points
 * (include/ear/dsp/gain_interpolator.hpp:42-43) without leaving HBM.  sample_idx stays a host
 * array (it drives the host-side segmentation).  Returns after `stream` has drained. */
int earb200_interp_set_points_device(earb200_interp *interp, int64_t n_points,
                                     const int64_t *sample_idx, const float *d_gains, void *stream);
/* process(block_start, nsamples, in, out): in is n_in planar channels of nsamples
 * floats, channel c at in + c*in_stride; out likewise (overwritten). */
int earb200_interp_process_host(earb200_interp *interp, int64_t block_start,
                                int64_t nsamples, const float *in, int64_t in_stride,
                                float *out, int64_t out_stride);
/* device buffers, asynchronous on `stream`.  d_out must not alias d_in (the kernels read the
 * input through the read-only path while other CTAs already store output). */
int earb200_interp_process_device(earb200_interp *interp, int64_t block_start,
                                  int64_t nsamples, const float *d_in, int64_t in_stride,
                                  float *d_out, int64_t out_stride, void *stream);
/* extension (not in libear): with accumulate on, process ADDS its result to `out` instead of
 * overwriting it, so several interpolators (e.g. groups of up to 256 objects each) mix into one
 * loudspeaker bus without an extra pass. */
int earb200_interp_set_accumulate(earb200_interp *interp, int enable);
int64_t earb200_interp_launch_count(const earb200_interp *interp);
int earb200_interp_set_profiling(earb200_interp *interp, int enable);
int earb200_interp_kernel_time(earb200_interp *interp, double *ms, int64_t *launches);

/* ------------------------------------------------- node-wide loudspeaker bus (exchange step)
 * Object-sharded mixing: libear's contract is "gains are applied to this channel, and summed with
 * other objects into an n-channel bus" (include/ear/gain_calculators.hpp:43-56, docs/dsp.rst:40-71);
 * with the objects spread over the GPUs of one node the sum over GPUs is the one exchange step of
 * the path.  The bus [n_channels][n_samples] is cut into one time slice per rank.  A rank renders
 * its objects with earb200_interp_process_device_bus: the render kernel's epilogue stores every
 * tile straight into the staging slot the rank has at the slice's owner (peer stores over NVLink,
 * riding on the render), and earb200_bus_reduce finishes the step: flags published at system
 * scope, the owner's sum over the slots in rank order (deterministic), optionally stored into every
 * rank's full bus.  No collective library is involved; the host program only has to hand one
 * 64-byte IPC handle per rank around (or, with all ranks in one process, call connect_local). */
typedef struct earb200_bus earb200_bus;

/* at most 8 ranks (the GPUs of one node); n_samples = length of the bus in samples */
int earb200_bus_create(int device, int rank, int n_ranks, int n_channels, int64_t n_samples,
                       earb200_bus **out);
void earb200_bus_destroy(earb200_bus *bus);
/* one process per GPU: every rank exports its handle (earb200_bus_ipc_handle_bytes() bytes), the
 * host program gathers them in rank order and every rank opens the array */
int earb200_bus_ipc_handle_bytes(void);
int earb200_bus_get_ipc_handle(earb200_bus *bus, void *handle_out);
int earb200_bus_open_ipc(earb200_bus *bus, const void *handles_in_rank_order);
/* all ranks in this process: the handles of all ranks, in rank order */
int earb200_bus_connect_local(earb200_bus *const *buses, int n);
/* the slice [t0, t1) of the bus that `rank` owns */
int earb200_bus_slice(const earb200_bus *bus, int rank, int64_t *t0, int64_t *t1);
/* GainInterpolator::process of this rank's objects into the bus samples [bus_t0, bus_t0 + nsamples).
 * d_partial (row pitch partial_stride) is this rank's private partial bus: needed in accumulate
 * mode (groups of objects mixed one after the other: all but the last call go through
 * earb200_interp_process_device into d_partial, the last one through this call, which reads
 * d_partial and stores the sum to the bus), for more than 128 input channels and for calls that are
 * not whole aligned tiles (those are rendered into d_partial and pushed); otherwise it may be NULL. */
int earb200_interp_process_device_bus(earb200_interp *interp, int64_t block_start, int64_t nsamples,
                                      const float *d_in, int64_t in_stride, float *d_partial,
                                      int64_t partial_stride, earb200_bus *bus, int64_t bus_t0,
                                      void *stream);
/* a partial bus rendered some other way: [n_channels][n] for the bus samples [t0, t0 + n) */
int earb200_bus_push(earb200_bus *bus, const float *d_partial, int64_t partial_stride, int64_t t0,
                     int64_t n, void *stream);
/* Tells every peer that this rank's tiles of the epoch are in place (stream order: after the
 * renders queued so far).  earb200_bus_reduce does it itself when it has not been done; the separate
 * call is for hosts that drive several ranks through ONE stream or device, where every rank has to
 * publish before any rank starts waiting. */
int earb200_bus_publish(earb200_bus *bus, void *stream);
/* Ends the epoch (once per rank, after the rank's renders of the epoch, on the same stream): waits
 * for the tiles of all ranks and sums the slots of this rank's slice in rank order.
 * d_out (nullable): this rank's slice of the summed bus, [n_channels][t1 - t0], row pitch
 * out_stride (reduce-scatter).  gather != 0: the summed slice also goes to every rank's full bus
 * (all-reduce); queue earb200_bus_wait_result before reading earb200_bus_result(), and consume the
 * result on the same stream before the next epoch's reduce. */
int earb200_bus_reduce(earb200_bus *bus, float *d_out, int64_t out_stride, int gather, void *stream);
int earb200_bus_wait_result(earb200_bus *bus, void *stream);
const float *earb200_bus_result(const earb200_bus *bus);   /* [n_channels][...], device memory */
int64_t earb200_bus_result_stride(const earb200_bus *bus); /* its row pitch in floats */
/* the samples [t0, t0 + n) of the full bus into a planar device buffer (row pitch dst_stride) */
int earb200_bus_copy_result(const earb200_bus *bus, float *d_dst, int64_t dst_stride, int64_t t0,
                            int64_t n, void *stream);
int64_t earb200_bus_launch_count(const earb200_bus *bus);

/* ----------------------------------------------------- block convolver (K4)
 * replaces ear::dsp::block_convolver::{Context,Filter,BlockConvolver}
 * (include/ear/dsp/block_convolver.hpp:28-112, src/dsp/block_convolver_impl.cpp) and
 * the FFT plug-in it is built on (include/ear/fft.hpp:14-67): a bank of n_channels
 * independent uniformly-partitioned overlap-add convolvers of block size B. */
typedef struct earb200_convolver earb200_convolver;

/* block_size must be a power of two in [16, 4096]; num_blocks = max partitions */
int earb200_convolver_create(int64_t block_size, int n_channels, int64_t num_blocks,
                             int device, earb200_convolver **out);
void earb200_convolver_destroy(earb200_convolver *conv);
/* Filter::Filter(ctx, n, taps): returns the filter id and its number of blocks */
int earb200_convolver_add_filter(earb200_convolver *conv, int64_t n_taps,
                                 const float *taps, int *filter_id, int *num_blocks);
/* set_filter / crossfade_filter on one channel (channel < 0: all channels);
 * filter_id < 0 = no filter (unset_filter / fade_down) */
int earb200_convolver_set_filter(earb200_convolver *conv, int channel, int filter_id);
int earb200_convolver_crossfade_filter(earb200_convolver *conv, int channel, int filter_id);
/* process n_blocks consecutive blocks per channel; channel c at in + c*in_stride.
 * in may be NULL (silence, as process(nullptr, out)). */
int earb200_convolver_process_host(earb200_convolver *conv, int64_t n_blocks,
                                   const float *in, int64_t in_stride, float *out,
                                   int64_t out_stride);
/* device buffers, asynchronous on `stream`; in-place use (d_out == d_in) is NOT supported: a
 * run of blocks re-reads the input block before its first one (halo) after other CTAs may have
 * stored their output. */
int earb200_convolver_process_device(earb200_convolver *conv, int64_t n_blocks,
                                     const float *d_in, int64_t in_stride, float *d_out,
                                     int64_t out_stride, void *stream);
int64_t earb200_convolver_launch_count(const earb200_convolver *conv);
int earb200_convolver_set_profiling(earb200_convolver *conv, int enable);
int earb200_convolver_kernel_time(earb200_convolver *conv, double *ms, int64_t *launches);

/* ------------------------------------------------------------ FFT plug-in (a9)
 * replaces ear::FFTImpl<float>::plan(n_fft) / FFTPlan::transform_forward / transform_reverse
 * (include/ear/fft.hpp:27-62; KISS implementation src/fft_kiss.cpp:52-99) as a BATCHED
 * transform: `batch` independent n_fft-point real transforms per call.  forward: [batch][n_fft]
 * reals -> [batch][n_fft/2+1] complex (interleaved re, im; no packing); reverse: the inverse,
 * unnormalised (forward then reverse scales by n_fft), imaginary parts of bins 0 and n_fft/2
 * ignored.  The block convolver runs the same butterflies inside its own kernel. */
typedef struct earb200_fft earb200_fft;

int earb200_fft_create(int64_t n_fft, int device, earb200_fft **out); /* power of two, 32..8192 */
void earb200_fft_destroy(earb200_fft *fft);
int64_t earb200_fft_size(const earb200_fft *fft);
int earb200_fft_forward_host(earb200_fft *fft, int64_t batch, const float *in, float *out);
int earb200_fft_reverse_host(earb200_fft *fft, int64_t batch, const float *in, float *out);
int earb200_fft_forward_device(earb200_fft *fft, int64_t batch, const float *d_in, float *d_out,
                               void *stream);
int earb200_fft_reverse_device(earb200_fft *fft, int64_t batch, const float *d_in, float *d_out,
                               void *stream);
int64_t earb200_fft_launch_count(const earb200_fft *fft);

/* -------------------------------------------------------------- delay line
 * replaces ear::dsp::DelayBuffer (include/ear/dsp/delay_buffer.hpp:12-29,
 * src/dsp/delay_buffer_impl.cpp:12-47): n_channels planar channels delayed by `delay`
 * samples; the delay memory persists on the device between calls. */
typedef struct earb200_delay earb200_delay;

int earb200_delay_create(int n_channels, int64_t delay, int device, earb200_delay **out);
void earb200_delay_destroy(earb200_delay *d);
int64_t earb200_delay_get_delay(const earb200_delay *d);
/* channel c at in + c*in_stride / out + c*out_stride; in may be NULL (silence) */
int earb200_delay_process_host(earb200_delay *d, int64_t nsamples, const float *in,
                               int64_t in_stride, float *out, int64_t out_stride);
/* device buffers, asynchronous on `stream`; in-place use (d_out == d_in) is NOT supported:
 * out[t] reads in[t - delay], which another thread may already have overwritten. */
int earb200_delay_process_device(earb200_delay *d, int64_t nsamples, const float *d_in,
                                 int64_t in_stride, float *d_out, int64_t out_stride,
                                 void *stream);
/* extension: out = delay(in) + add in one pass (direct bus + decorrelated diffuse bus);
 * d_add must not alias d_out unless it is the same pointer with the same stride */
int earb200_delay_process_add_device(earb200_delay *d, int64_t nsamples, const float *d_in,
                                     int64_t in_stride, const float *d_add, int64_t add_stride,
                                     float *d_out, int64_t out_stride, void *stream);
/* extension: out = BlockConvolver(in) + DelayBuffer(direct) in one pass (the decorrelated diffuse bus
 * plus the compensation-delayed direct bus of the Objects signal flow, docs/dsp.rst:40-71); both
 * handles advance as if their process() had been called.  d_out must alias neither input. */
int earb200_convolver_process_add_delayed_device(earb200_convolver *conv, earb200_delay *d,
                                                 int64_t n_blocks, const float *d_in, int64_t in_stride,
                                                 const float *d_direct, int64_t direct_stride,
                                                 float *d_out, int64_t out_stride, void *stream);
int64_t earb200_delay_launch_count(const earb200_delay *d);

/* ------------------------------------------------ Objects signal flow (8f.1)
 * The rendering chain libear's documentation prescribes for typeDefinition == "Objects"
 * (docs/dsp.rst:40-71, include/ear/gain_calculators.hpp:43-56) as one device pipeline:
 * direct and diffuse buses by two gain interpolators, the diffuse bus through the per-channel
 * decorrelators (designDecorrelators, block size 512), the direct bus through the compensation
 * delay (255 samples), summed.  Calls process whole blocks of 512 samples. */
typedef struct earb200_objects_renderer earb200_objects_renderer;

int earb200_objects_renderer_create(const earb200_layout *layout, int n_objects, int device,
                                    earb200_objects_renderer **out);
void earb200_objects_renderer_destroy(earb200_objects_renderer *r);
int earb200_objects_renderer_num_outputs(const earb200_objects_renderer *r);
int64_t earb200_objects_renderer_block_size(const earb200_objects_renderer *r); /* 512 */
int earb200_objects_renderer_delay(const earb200_objects_renderer *r);          /* 255 */
int64_t earb200_objects_renderer_launch_count(const earb200_objects_renderer *r);
/* interpolation points: direct / diffuse gains [n_points][n_objects][C_out] (e.g. the output of
 * earb200_objects_calculate_* per metadata block), host buffers */
int earb200_objects_renderer_set_gains(earb200_objects_renderer *r, int64_t n_points,
                                       const int64_t *sample_idx, const float *direct,
                                       const float *diffuse);
/* the same with device-resident gains (earb200_objects_calculate_device output): no host round trip */
int earb200_objects_renderer_set_gains_device(earb200_objects_renderer *r, int64_t n_points,
                                              const int64_t *sample_idx, const float *d_direct,
                                              const float *d_diffuse, void *stream);
/* n_blocks * 512 samples per channel: in = n_objects planar channels, out = C_out feeds */
int earb200_objects_renderer_process_host(earb200_objects_renderer *r, int64_t block_start,
                                          int64_t n_blocks, const float *in, int64_t in_stride,
                                          float *out, int64_t out_stride);
int earb200_objects_renderer_process_device(earb200_objects_renderer *r, int64_t block_start,
                                            int64_t n_blocks, const float *d_in, int64_t in_stride,
                                            float *d_out, int64_t out_stride, void *stream);

/* ------------------------------- Cartesian <-> polar metadata conversion (8f.3)
 * replaces ear::conversion::{pointCartToPolar, pointPolarToCart, extentCartToPolar,
 * extentPolarToCart} and, through the facade, toPolar / toCartesian
 * (include/ear/conversion.hpp:14-71, src/conversion.cpp:127-263) for a batch of objects:
 * a pre-pass that turns Cartesian-authored Objects blocks into the polar blocks
 * earb200_objects_calculate_* accepts.  pos [n][3] is X, Y, Z on the Cartesian side and
 * azimuth, elevation, distance on the polar side; extent [n][3] is width, height, depth
 * (nullable together with extent_out: positions only).  A position libear would throw
 * ear::internal_error for (no sector found / p out of range: non-finite input) gives NaN
 * outputs; the _host variants then return EARB200_INTERNAL_ERROR, the asynchronous _device
 * variants report it at the next earb200_conversion_check (which synchronises). */
typedef struct earb200_conversion earb200_conversion;

int earb200_conversion_create(int device, earb200_conversion **out);
void earb200_conversion_destroy(earb200_conversion *conv);
int earb200_conversion_cart_to_polar_host(earb200_conversion *conv, int64_t n, const double *pos,
                                          const double *extent, double *pos_out,
                                          double *extent_out);
int earb200_conversion_polar_to_cart_host(earb200_conversion *conv, int64_t n, const double *pos,
                                          const double *extent, double *pos_out,
                                          double *extent_out);
int earb200_conversion_cart_to_polar_device(earb200_conversion *conv, int64_t n,
                                            const double *d_pos, const double *d_extent,
                                            double *d_pos_out, double *d_extent_out, void *stream);
int earb200_conversion_polar_to_cart_device(earb200_conversion *conv, int64_t n,
                                            const double *d_pos, const double *d_extent,
                                            double *d_pos_out, double *d_extent_out, void *stream);
int earb200_conversion_check(earb200_conversion *conv, void *stream);
int64_t earb200_conversion_launch_count(const earb200_conversion *conv);

/* ------------------------------------------------ HOA decoder design (K5)
 * replaces ear::GainCalculatorHOA (include/ear/gain_calculators.hpp:67-82,
 * src/hoa/gain_calculator_hoa.cpp:8-72, src/hoa/hoa.hpp:88-143). */
typedef struct earb200_hoa earb200_hoa;

enum {
  EARB200_WARN_HOA_SCREENREF = 4,  /* Warning::Code::HOA_SCREENREF_NOT_IMPLEMENTED  */
  EARB200_WARN_HOA_NFCREFDIST = 5  /* Warning::Code::HOA_NFCREFDIST_NOT_IMPLEMENTED */
};

int earb200_hoa_create(const earb200_layout *layout, int device, earb200_hoa **out);
void earb200_hoa_destroy(earb200_hoa *hoa);
/* decode [n_channels][C_out] row-major double (one vector of per-loudspeaker gains
 * per HOA input channel).  warnings: bitmask of (1 << code) raised, nullable. */
int earb200_hoa_calculate(earb200_hoa *hoa, int n_channels, const int *orders,
                          const int *degrees, const char *normalization,
                          double nfc_ref_dist, int screen_ref, double *decode,
                          uint32_t *warnings);

/* ------------------------------------- DirectSpeakers gain calculator (a12)
 * replaces ear::GainCalculatorDirectSpeakers (include/ear/gain_calculators.hpp:18-34,
 * src/direct_speakers/gain_calculator_direct_speakers.cpp:57-320).  The routing decision
 * is host work (labels, ITU-R BS.2127 mapping rules, bounds); channels that fall through
 * to the point-source panner are panned by K1 in one launch per call. */
typedef struct earb200_direct_speakers earb200_direct_speakers;

/* flat form of ear::DirectSpeakersTypeMetadata (include/ear/metadata.hpp:22-72) */
typedef struct earb200_direct_speakers_metadata {
  const char *const *speaker_labels; /* speakerLabels, in AXML order */
  int32_t n_speaker_labels;
  int32_t cartesian_position; /* position holds a CartesianSpeakerPosition (libear: not_implemented) */
  double position[3];         /* azimuth, elevation, distance */
  double bounds[6];           /* azimuthMin, azimuthMax, elevationMin, elevationMax, distanceMin, distanceMax */
  uint32_t has_bounds;        /* bit i set: bounds[i] is present */
  uint32_t screen_edge_lock;  /* bit 0: horizontal, bit 1: vertical (libear: not_implemented) */
  double low_pass, high_pass; /* channelFrequency */
  uint32_t has_frequency;     /* bit 0: lowPass present, bit 1: highPass present */
  uint32_t reserved;
  const char *audio_pack_format_id; /* nullable, e.g. "AP_00010003" */
} earb200_direct_speakers_metadata;

enum {
  EARB200_WARN_FREQ_SPEAKERLABEL_LFE_MISMATCH = 1, /* Warning::Code (include/ear/warnings.hpp:13-17) */
  EARB200_WARN_FREQ_NOT_LFE = 2
};

/* additionalSubstitutions: n pairs from[i] -> to[i] (nullable when n == 0) */
int earb200_direct_speakers_create(const earb200_layout *layout, int device, int n_substitutions,
                                   const char *const *from, const char *const *to,
                                   earb200_direct_speakers **out);
void earb200_direct_speakers_destroy(earb200_direct_speakers *calc);
int earb200_direct_speakers_num_outputs(const earb200_direct_speakers *calc);
/* n channels; gains [n][C_out] row-major double; warnings [n] bitmask of (1 << code), nullable.
 * Host buffers. */
int earb200_direct_speakers_calculate(earb200_direct_speakers *calc, int64_t n,
                                      const earb200_direct_speakers_metadata *md, double *gains,
                                      uint32_t *warnings);
int64_t earb200_direct_speakers_launch_count(const earb200_direct_speakers *calc);

/* ---------------------------------------------------------- decorrelators
 * replaces ear::designDecorrelators / decorrelatorCompensationDelay
 * (include/ear/decorrelate.hpp:17-34, src/decorrelate.cpp:31-97).  Setup-time, host. */
int earb200_decorrelator_size(void);               /* 512 */
int earb200_decorrelator_compensation_delay(void); /* 255 */
/* filters [n_channels][512]; float_out / double_out nullable */
int earb200_design_decorrelators(const earb200_layout *layout, float *float_out,
                                 double *double_out);

#ifdef __cplusplus
}
#endif
#endif /* EARB200_H */

// POD tables shared between the host-side setup code and the device kernels.
//
// They are the flattened, device-friendly form of what libear builds as a tree of
// heap objects: the region list of configureFullPolarPanner
// (ref src/common/point_source_panner.cpp:473-557), the 1,652-point extent grid with
// its cached point-source gains (ref src/object_based/polar_extent.cpp:98-165) and the
// LFE scatter map of GainCalculatorObjectsImpl (ref gain_calculator_objects.cpp:51-52).
#pragma once
#include <stdint.h>

#define EARB_MAX_SPK 24        // S: real non-LFE loudspeakers (9+10+3 has 22)
#define EARB_MAX_OUT 32        // C_out: all output channels including LFE
#define EARB_MAX_AUG 32        // A: real + "extra" vertical copies (9+10+3 has 27)
#define EARB_MAX_REGIONS 48    // top-level regions (9+10+3 has 37)
#define EARB_MAX_TRI 64        // triplets incl. the inner triplets of virtual n-gons
#define EARB_MAX_QUAD 16
#define EARB_MAX_NGON 2
#define EARB_MAX_NGON_VERTS 16

enum { EARB_REGION_TRIPLET = 0, EARB_REGION_QUAD = 1, EARB_REGION_NGON = 2 };

struct PspTriplet {
  // inverse of the 3x3 matrix whose rows are the speaker positions; gains are
  // pv[j] = sum_i pos[i] * basis[3*i + j]            (ref point_source_panner.cpp:36-50)
  double basis[9];
  int32_t ch[3];  // output channel of each gain (augmented index, or n-gon local index)
  int32_t pad;
};

struct PspQuad {
  double polyX[9];  // row-major; poly = polyX * pos           (ref :132-145, :174-186)
  double polyY[9];
  double pos[12];   // speaker positions [4][3] in facet order  (ref :125)
  int32_t order[4]; // ngonVertexOrder of the four vertices     (ref :104, :123)
  int32_t ch[4];    // augmented channel of each vertex
};

struct PspNgon {
  int32_t n;          // number of real vertices
  int32_t first_tri;  // index of the first of n inner triplets in PspTables::tri
  int32_t ch[EARB_MAX_NGON_VERTS];  // augmented channel of each vertex
  double centre_downmix;            // 1/sqrt(n)                (ref :512-514)
};

struct PspRegion {
  int32_t kind;
  int32_t index;
};

struct PspTables {
  int32_t n_regions;
  int32_t n_aug;    // A (without the virtual vertices)
  int32_t n_real;   // S: number of channels the panner returns
  int32_t stereo;   // 1: 0+2+0 -- tables are those of 0+5+0, followed by the stereo downmix
  PspRegion regions[EARB_MAX_REGIONS];
  PspTriplet tri[EARB_MAX_TRI];
  PspQuad quad[EARB_MAX_QUAD];
  PspNgon ngon[EARB_MAX_NGON];
  int32_t parent[EARB_MAX_AUG];  // augmented channel -> real channel (0/1 downmix matrix)
};

// extent grid geometry; the arrays themselves live in separate device buffers
#define EARB_EXT_BLOCK 32  // points per weight-evaluation block (one warp)

struct ExtentGrid {
  int32_t n_points;      // P (1,652 for the reference's nRows = 37 grid)
  int32_t n_points_pad;  // rounded up to a multiple of EARB_EXT_BLOCK
  int32_t n_blocks;      // n_points_pad / EARB_EXT_BLOCK
  int32_t n_spk;         // S
};

// one spreading pass of one object: the float context of
// PolarExtentCoreContext (ref src/object_based/polar_extent_core.hpp:12-42)
struct __attribute__((aligned(16))) ExtentPass {
  float basis[9];
  float circle_test[2];
  float circle_centre[2];
  float cos_start, cos_end, sin_start, sin_end;
  float m, c;
  int32_t is_circular;
  int32_t active;  // 0: spread branch skipped (amount_spread <= 1e-10)
  int32_t slot;    // 2 * object + pass: where the core kernel writes this pass's results
  int32_t pad[2];
};  // 24 words = 96 bytes (loaded as six 16-byte vectors)

// per-object scalars carried from the prelude to the epilogue
struct ObjectAux {
  double amount_spread[2];  // per pass
  int32_t n_passes;         // 1 or 2 (2 when depth != 0)
  int32_t has_point;        // bit p set: pass p needs the point-source gains;
                            // bit 2 + p set: pass p has a spread part (an ExtentPass exists)
};

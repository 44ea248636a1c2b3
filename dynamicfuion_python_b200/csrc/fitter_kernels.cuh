// Kernel launchers of the fused per-iteration path used by the DeformableMeshToImageFitter host class (fitter.cu).
//
// Device data layout (all 16-byte aligned, resident in the fitter workspace for the whole fit):
//   canonical / warped vertices and normals   float4 [V]            (xyz, pad)      128-bit coalesced loads
//   triangles                                 int4   [F]            (v0, v1, v2, pad)
//   vertex anchors / weights                  int32 / f32 [V, A]
//   vertex-anchor Jacobian records            float4 [V, A, 2]      (-wR(v-g), w) , (-wRn, node id bits)
//   face-anchor table                         int2   [F, 3A] + u8 count [F]   (node, packed per-vertex anchor slots)
//   pixel keys                                u64    [H*W]          (depth bits << 32 | face), ~0 = empty
//   reference points                          float4 [H*W]          (xyz, mask)
//   node accumulators                         2 x u64 [N, 32]       21 upper-triangular J^T J + 6 J^T r + pad, 103-bit fixed point (fixed_accum.cuh)
#pragma once
#include "host_util.h"
#include "raster_core.cuh"

namespace nnrt_b200 {

constexpr int kAccumStride = 32;
constexpr int kMaxStemDegree = 8;

struct DataTermParams {
	int height, width;
	Intrinsics pinhole;   // camera intrinsics as given (unprojection)
	Intrinsics ndc;       // NDC intrinsics (projection / Jacobians)
	float max_depth;
	int perspective_correct;
	int use_tukey;
	float tukey_cutoff;
	float blur_radius;    // NDC
	int row_begin, row_end;  // image rows [row_begin, row_end) this GPU rasterises and assembles (whole image unless pixel-sharded)
	int write_diagnostics;   // per-pixel face / residual / mask of this iteration are kept (the last iteration of a fit; every split or profiled one)
};

struct FitterDeviceData {
	// mesh
	const float4* positions; const float4* normals; const int4* triangles; int64_t vertex_count; int64_t face_count;
	const int32_t* anchors; const float* anchor_weights; int anchor_count;
	// graph
	const float* nodes; int node_count; int arrow_base;
	// per-iteration buffers
	float4* warped_positions; float4* warped_normals; float4* vertex_jacobians;
	unsigned long long* keys; BandItem* band_queue; int* counters;  // counters[0] = queue count, [1] = overflow flag, [2] = not-PD flag, [3] = accumulator range flag
	int band_queue_capacity;
	const float4* reference_points;
	unsigned long long* accumulators;  // [N, kAccumStride, 2] fixed-point words (fixed_accum.cuh)
	int32_t* diag_pixel_faces; float* diag_residuals; uint8_t* diag_mask;  // diagnostics of the last iteration
};

void launch_warp_vertices(const FitterDeviceData& d, const float* rotations, const float* translations, const Rigid& extrinsics, cudaStream_t s);
void launch_face_raster(const FitterDeviceData& d, const DataTermParams& p, const NdcCamera& cam, cudaStream_t s);
void launch_band_raster(const FitterDeviceData& d, const DataTermParams& p, const NdcCamera& cam, cudaStream_t s);
int launch_assemble(const FitterDeviceData& d, const DataTermParams& p, cudaStream_t s);
void launch_fill_keys(unsigned long long* keys, int64_t n, cudaStream_t s);

}  // namespace nnrt_b200

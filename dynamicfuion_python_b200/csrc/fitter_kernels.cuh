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
constexpr int kMaxPeers = 16;

// Pixel-sharded fits without a library collective: every GPU's accumulators live in a buffer its peers can read (NVLink peer memory),
// and the kernel that consumes the sums (system_stem_kernel) gathers and adds the ranks' words itself. Two accumulator buffers
// alternate by iteration parity so that a GPU one iteration ahead never writes what a peer is still reading; `flags[r]` on every GPU
// holds 1 + the last iteration rank r has finished assembling. See fitter.cu::system_stem_kernel for the protocol.
struct PeerReduce {
	int world, rank;                                  // world <= 1: off
	long long parity_words;                           // 64-bit words per accumulator buffer = N * kAccumStride * 2
	unsigned long long* epoch;                        // this GPU's iteration counter (device memory; the kernels read and advance it)
	unsigned int* done;                               // warps of system_stem_kernel that have read the epoch and the sums
	unsigned long long* flags;                        // this GPU's flag row [kMaxPeers]
	unsigned long long* peer_flags[kMaxPeers];        // every rank's flag row, as mapped into this process
	const unsigned long long* peer_accum[kMaxPeers];  // every rank's [2][parity_words] accumulators, as mapped into this process ([rank] = own)
};

// Work item of the fitter's band queue: rows [y0, y1] x columns [u0, u1] of one face, with the face's projected vertices. The
// stage-level rasteriser's 16-byte BandItem makes the draining warp re-derive the face (item -> triangle -> three vertices -> project
// -> pixel box: three dependent round trips per item, which is what bound band_raster_kernel at 3840 x 2160: 0.19 ms for ~100 k
// items); here the thread that queued the item already had all of it in registers, and the drain reads four 16-byte words.
struct __align__(16) FaceBandItem {
	int face, y0, y1, u0;
	int u1; float z0, z1, z2;
	float2 v0, v1;
	float2 v2; float pad0, pad1;
};
static_assert(sizeof(FaceBandItem) == 64, "four 16-byte words per queued band");

struct DataTermParams {
	int height, width;
	Intrinsics pinhole;   // camera intrinsics as given (unprojection)
	float pinhole_inverse_fx, pinhole_inverse_fy;  // 1.0f / fx, 1.0f / fy, correctly rounded on the host (common.cuh::SharedDivisor)
	Intrinsics ndc;       // NDC intrinsics (projection / Jacobians)
	float max_depth;
	int perspective_correct;
	int use_tukey;
	float tukey_cutoff;
	float blur_radius;    // NDC
	int row_begin, row_end;  // image rows [row_begin, row_end) this GPU rasterises and assembles (whole image unless pixel-sharded)
	// pixel-sharded fits, balanced form: the rows are cut into stripes of `stripe_rows` (a multiple of the 8-row assembly tile) dealt
	// round-robin to `stripe_count` GPUs; this GPU owns stripes stripe_index, stripe_index + stripe_count, ... (8, 0, 1 = every row)
	int stripe_rows, stripe_index, stripe_count;
	const float* ndc_x; const float* ndc_y;  // pixel centres in NDC, [width] and [height] (raster_core.cuh::RasterParams); filled at fitter_create
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
	unsigned long long* keys; FaceBandItem* band_queue; int* counters;  // counters[0] = queue count, [1] = overflow flag, [2] = not-PD flag, [3] = accumulator range flag
	int band_queue_capacity;
	const float4* reference_points;
	unsigned long long* accumulators;  // [N, kAccumStride, 2] fixed-point words (fixed_accum.cuh)
	const unsigned long long* accum_epoch; long long accum_parity_words;  // PeerReduce: this iteration's buffer is accumulators + (*accum_epoch & 1) * accum_parity_words
	int32_t* diag_pixel_faces; float* diag_residuals; uint8_t* diag_mask;  // diagnostics of the last iteration
};

__host__ __device__ __forceinline__ bool row_owned(const DataTermParams& p, int v) {
	return p.stripe_count <= 1 || ((v - p.row_begin) / p.stripe_rows) % p.stripe_count == p.stripe_index;
}

void launch_warp_vertices(const FitterDeviceData& d, const float* rotations, const float* translations, const Rigid& extrinsics, cudaStream_t s);
void launch_face_raster(const FitterDeviceData& d, const DataTermParams& p, const NdcCamera& cam, cudaStream_t s);
void launch_band_raster(const FitterDeviceData& d, const DataTermParams& p, const NdcCamera& cam, cudaStream_t s);
int launch_assemble(const FitterDeviceData& d, const DataTermParams& p, cudaStream_t s);
void launch_fill_keys(unsigned long long* keys, int64_t n, cudaStream_t s);
void launch_fill_pixel_centres(float* ndc_x, float* ndc_y, int height, int width, cudaStream_t s);

}  // namespace nnrt_b200

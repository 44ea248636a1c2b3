// Mesh and ordered-point-cloud normals (nnrt.geometry.functional.compute_{triangle,vertex,ordered_point_cloud}_normals): the
// canonical mesh handed to the fitter must carry vertex normals (DeformableMeshToImageFitter.cpp:346-348).
//
// The reference accumulates triangle normals into vertex normals with float atomics, so its CUDA result depends on scheduling. Here
// the corners incident to a vertex are gathered through a CSR built on the device (count, scan, fill, rank by face index) and summed
// in ascending face order: reproducible, and bit-identical to the reference's single-thread CPU result (the oracle).
// Arithmetic follows Eigen: cross = (a.y b.z - a.z b.y, ...), squaredNorm = x + (y + z), normalize divides by sqrt(z) when z > 0.
#include "host_util.h"
#include "device_scan.cuh"
#include "../../include/nnrt_b200.h"

namespace nnrt_b200 {
namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ float3 cross3(float3 a, float3 b) {
	return make_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
// Eigen MatrixBase::normalize() / normalized() (Eigen/src/Core/Dot.h:119-148)
__device__ __forceinline__ float3 normalized3(float3 v) {
	const float z = sum3(v.x * v.x, v.y * v.y, v.z * v.z);
	if (z > 0.f) {
		const float norm = sqrtf(z);
		return make_float3(__fdiv_rn(v.x, norm), __fdiv_rn(v.y, norm), __fdiv_rn(v.z, norm));
	}
	return v;
}
// kernel::mesh::NormalizeVectors3d (geometry/functional/kernel/NormalsOperationsImpl.h:72-93): NaN -> (0, 0, 1)
__device__ __forceinline__ float3 normalize_or_up(float3 v) {
	v = normalized3(v);
	if (isnan(v.x)) v = make_float3(0.f, 0.f, 1.f);
	return v;
}
__device__ __forceinline__ float3 load3(const float* __restrict__ p, int64_t i) { return make_float3(p[i * 3], p[i * 3 + 1], p[i * 3 + 2]); }
__device__ __forceinline__ void store3(float* __restrict__ p, int64_t i, float3 v) { p[i * 3] = v.x; p[i * 3 + 1] = v.y; p[i * 3 + 2] = v.z; }

// geometry/functional/kernel/NormalsOperationsImpl.h:38-69 (+ NormalizeVectors3d when `normalized`)
__global__ void __launch_bounds__(kThreads) triangle_normals_kernel(const float* __restrict__ positions, const int64_t* __restrict__ triangles, int64_t face_count,
                                                                    bool normalized, float* __restrict__ normals) {
	const int64_t f = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (f >= face_count) return;
	const float3 v0 = load3(positions, triangles[f * 3]), v1 = load3(positions, triangles[f * 3 + 1]), v2 = load3(positions, triangles[f * 3 + 2]);
	float3 n = cross3(v1 - v0, v2 - v0);
	if (normalized) n = normalize_or_up(n);
	store3(normals, f, n);
}

__global__ void __launch_bounds__(kThreads) count_corners_kernel(const int64_t* __restrict__ triangles, int64_t corner_count, int* __restrict__ counts) {
	const int64_t c = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (c < corner_count) atomicAdd(counts + triangles[c], 1);
}
__global__ void __launch_bounds__(kThreads) fill_corners_kernel(const int64_t* __restrict__ triangles, int64_t corner_count, const int* __restrict__ starts,
                                                                int* __restrict__ fill, int* __restrict__ corners) {
	const int64_t c = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (c >= corner_count) return;
	const int64_t v = triangles[c];
	corners[starts[v] + atomicAdd(fill + v, 1)] = (int) c;
}
// geometry/functional/kernel/NormalsOperationsImpl.h:96-165: vertex normal = sum of the normals of the triangles that reference it,
// one addend per corner, here in ascending corner (= face) order. The arrival-ordered corner list is walked in rank order: each step
// takes the smallest corner index above the previous one (degree^2 / 2 comparisons; mesh degrees are ~6).
__global__ void __launch_bounds__(kThreads) vertex_normals_kernel(const int* __restrict__ starts, const int* __restrict__ corners,
                                                                  const float* __restrict__ triangle_normals, int64_t vertex_count, bool normalized,
                                                                  float* __restrict__ vertex_normals) {
	const int64_t v = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (v >= vertex_count) return;
	const int begin = starts[v], end = starts[v + 1];
	float3 sum = make_float3(0.f, 0.f, 0.f);
	int previous = -1;
	for (int step = begin; step < end; step++) {
		int next = 0x7fffffff;
		for (int k = begin; k < end; k++) {
			const int c = corners[k];
			if (c > previous && c < next) next = c;
		}
		previous = next;
		sum = sum + load3(triangle_normals, next / 3);
	}
	if (normalized) sum = normalize_or_up(sum);
	store3(vertex_normals, v, sum);
}

// geometry/functional/kernel/NormalsOperationsImpl.h:172-214
__global__ void __launch_bounds__(kThreads) ordered_point_cloud_normals_kernel(const float* __restrict__ points, int height, int width,
                                                                               float* __restrict__ normals) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= (int64_t) height * width) return;
	const int y = (int) (i / width), x = (int) (i % width);
	float3 n = make_float3(0.f, 0.f, 0.f);
	if (!(x == 0 || x == width - 1 || y == 0 || y == height - 1)) {
		const float3 d_horizontal = load3(points, i + 1) - load3(points, i - 1);
		const float3 d_vertical = load3(points, i - width) - load3(points, i + width);
		n = normalized3(cross3(d_horizontal, d_vertical));
		if (n.z > 0) n = make_float3(-n.x, -n.y, -n.z);
	}
	store3(normals, i, n);
}

__global__ void __launch_bounds__(kThreads) normalize_vectors_kernel(float* __restrict__ vectors, int64_t count) {
	const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < count) store3(vectors, i, normalize_or_up(load3(vectors, i)));
}

}  // namespace
}  // namespace nnrt_b200

using namespace nnrt_b200;

extern "C" int nnrt_b200_compute_triangle_normals(const float* vertex_positions, int64_t vertex_count, const int64_t* triangles, int64_t face_count,
                                                  int32_t normalized, float* triangle_normals, void* stream) {
	if (face_count == 0) return kOk;
	NNRT_CHECK_ARG(vertex_positions && triangles && triangle_normals, "compute_triangle_normals: null pointer argument");
	NNRT_CHECK_ARG(vertex_count > 0 && face_count > 0, "compute_triangle_normals: vertex_count and face_count must be positive");
	triangle_normals_kernel<<<blocks_for(face_count, kThreads), kThreads, 0, (cudaStream_t) stream>>>(vertex_positions, triangles, face_count, normalized != 0,
	                                                                                                triangle_normals);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_compute_vertex_normals(const int64_t* triangles, int64_t face_count, const float* triangle_normals, int64_t vertex_count,
                                                int32_t normalized, float* vertex_normals, void* stream) {
	if (vertex_count == 0) return kOk;
	NNRT_CHECK_ARG(vertex_normals && vertex_count > 0 && vertex_count < (1ll << 31) - 2, "compute_vertex_normals: bad vertex_normals / vertex_count");
	NNRT_CHECK_ARG(face_count == 0 || (triangles && triangle_normals), "compute_vertex_normals: null pointer argument");
	NNRT_CHECK_ARG(face_count >= 0 && face_count * 3 < (1ll << 31), "compute_vertex_normals: face_count out of range");
	cudaStream_t s = (cudaStream_t) stream;
	const int64_t corner_count = face_count * 3;
	auto align = [](size_t b) { return (b + 255) & ~(size_t) 255; };
	const size_t ints = (size_t) vertex_count + 2;
	const size_t total = 2 * align(sizeof(int) * ints) + align(sizeof(int) * (size_t) (corner_count + 1)) + align(sizeof(int) * (size_t) scan_scratch_ints(vertex_count + 1));
	char* buffer = nullptr;
	Scratch scratch_guard(s);
	NNRT_CUDA(scratch_guard.alloc(&buffer, total));
	int* counts = (int*) buffer;                                   // becomes the CSR starts [V + 1]
	int* fill = (int*) (buffer + align(sizeof(int) * ints));       // adjacent to counts: one memset clears both
	int* corners = (int*) (buffer + 2 * align(sizeof(int) * ints));
	int* scratch = (int*) (buffer + 2 * align(sizeof(int) * ints) + align(sizeof(int) * (size_t) (corner_count + 1)));
	int rc = kOk;
	do {
		if (cudaMemsetAsync(counts, 0, align(sizeof(int) * ints) + sizeof(int) * ints, s) != cudaSuccess) { rc = kCudaError; break; }
		if (corner_count > 0) count_corners_kernel<<<blocks_for(corner_count, kThreads), kThreads, 0, s>>>(triangles, corner_count, counts);
		rc = exclusive_scan(counts, counts, vertex_count, scratch, s);
		if (rc != kOk) break;
		if (corner_count > 0) fill_corners_kernel<<<blocks_for(corner_count, kThreads), kThreads, 0, s>>>(triangles, corner_count, counts, fill, corners);
		vertex_normals_kernel<<<blocks_for(vertex_count, kThreads), kThreads, 0, s>>>(counts, corners, triangle_normals, vertex_count, normalized != 0, vertex_normals);
		if (cudaGetLastError() != cudaSuccess) { set_last_error("compute_vertex_normals: kernel launch failed"); rc = kCudaError; }
	} while (false);
	return rc;
}

// ---------------------------------------------------------------------------------------------------------------------------
// Allocation-free variant of compute_vertex_normals (no counterpart in the reference; added for
// callers that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). The CSR scratch
// lives in a file-scope device array, so every GPU that loads the library holds kStaticScratchBytes (~4.2 MB) until exit.
// All calls on one device share the array and it keeps the previous call's contents: the counters are cleared at the start of each
// call, and two calls that overlap on the GPU (different streams, no ordering between them) race.
namespace nnrt_b200 {
namespace {
constexpr int64_t kStaticMaxVertices = 1 << 17, kStaticMaxFaces = 1 << 18;
constexpr size_t align256(size_t b) { return (b + 255) & ~(size_t) 255; }
constexpr size_t kStaticCounterBytes = align256(sizeof(int) * (size_t) (kStaticMaxVertices + 2));
constexpr size_t kStaticCornerBytes = align256(sizeof(int) * (size_t) (kStaticMaxFaces * 3 + 1));
constexpr size_t kStaticScanBytes = align256(sizeof(int) * (size_t) ((kStaticMaxVertices + 1 + kScanTile - 1) / kScanTile + 1));
constexpr size_t kStaticScratchBytes = 2 * kStaticCounterBytes + kStaticCornerBytes + kStaticScanBytes;

static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

// In host code `g_static_scratch` is only a stand-in for the symbol: the device address is looked up once per device (with that
// device current) and cached by ordinal.
int static_scratch_address(char** address) {
	static std::atomic<char*> cached[64];
	int device = 0;
	NNRT_CUDA(cudaGetDevice(&device));
	NNRT_CHECK_ARG(device >= 0 && device < 64, "compute_vertex_normals_static_scratch: device ordinal %d not supported", device);
	char* p = cached[device].load(std::memory_order_acquire);
	if (p == nullptr) {
		NNRT_CUDA(cudaGetSymbolAddress((void**) &p, g_static_scratch));
		cached[device].store(p, std::memory_order_release);
	}
	*address = p;
	return kOk;
}
}  // namespace
}  // namespace nnrt_b200

static_assert(kStaticMaxVertices == NNRT_B200_STATIC_SCRATCH_MAX_VERTICES && kStaticMaxFaces == NNRT_B200_STATIC_SCRATCH_MAX_FACES,
              "include/nnrt_b200.h states the limits of the static scratch");

extern "C" int nnrt_b200_compute_vertex_normals_static_scratch(const int64_t* triangles, int64_t face_count, const float* triangle_normals,
                                                               int64_t vertex_count, int32_t normalized, float* vertex_normals, void* stream) {
	if (vertex_count == 0) return kOk;
	NNRT_CHECK_ARG(vertex_normals && vertex_count > 0, "compute_vertex_normals_static_scratch: bad vertex_normals / vertex_count");
	NNRT_CHECK_ARG(face_count == 0 || (triangles && triangle_normals), "compute_vertex_normals_static_scratch: null pointer argument");
	NNRT_CHECK_ARG(face_count >= 0, "compute_vertex_normals_static_scratch: face_count out of range");
	NNRT_CHECK_ARG(vertex_count <= kStaticMaxVertices && face_count <= kStaticMaxFaces,
	               "compute_vertex_normals_static_scratch: %lld vertices / %lld faces exceed the static scratch limit of %lld vertices / %lld faces; "
	               "use compute_vertex_normals for larger meshes",
	               (long long) vertex_count, (long long) face_count, (long long) kStaticMaxVertices, (long long) kStaticMaxFaces);
	char* buffer = nullptr;
	const int rc = static_scratch_address(&buffer);
	if (rc != kOk) return rc;
	cudaStream_t s = (cudaStream_t) stream;
	const int64_t corner_count = face_count * 3;
	int* counts = (int*) buffer;                                   // becomes the CSR starts [V + 1]
	int* fill = (int*) (buffer + kStaticCounterBytes);             // adjacent to counts: one memset clears both
	int* corners = (int*) (buffer + 2 * kStaticCounterBytes);
	int* scan_tiles = (int*) (buffer + 2 * kStaticCounterBytes + kStaticCornerBytes);
	// the previous call's counters are still there: clear them (corners and scan tiles are fully rewritten before they are read)
	NNRT_CUDA(cudaMemsetAsync(counts, 0, kStaticCounterBytes + sizeof(int) * (size_t) (vertex_count + 2), s));
	if (corner_count > 0) count_corners_kernel<<<blocks_for(corner_count, kThreads), kThreads, 0, s>>>(triangles, corner_count, counts);
	const int scan_rc = exclusive_scan(counts, counts, vertex_count, scan_tiles, s);
	if (scan_rc != kOk) return scan_rc;
	if (corner_count > 0) fill_corners_kernel<<<blocks_for(corner_count, kThreads), kThreads, 0, s>>>(triangles, corner_count, counts, fill, corners);
	vertex_normals_kernel<<<blocks_for(vertex_count, kThreads), kThreads, 0, s>>>(counts, corners, triangle_normals, vertex_count, normalized != 0, vertex_normals);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_normalize_vectors_3d(float* vectors, int64_t count, void* stream) {
	if (count == 0) return kOk;
	NNRT_CHECK_ARG(vectors && count > 0, "normalize_vectors_3d: null pointer or negative count");
	normalize_vectors_kernel<<<blocks_for(count, kThreads), kThreads, 0, (cudaStream_t) stream>>>(vectors, count);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

extern "C" int nnrt_b200_compute_ordered_point_cloud_normals(const float* point_positions, int32_t image_height, int32_t image_width, float* normals,
                                                             void* stream) {
	NNRT_CHECK_ARG(point_positions && normals, "compute_ordered_point_cloud_normals: null pointer argument");
	NNRT_CHECK_ARG(image_height > 0 && image_width > 0, "Source image size must be positive, got %d x %d.", image_height, image_width);
	const int64_t count = (int64_t) image_height * image_width;
	ordered_point_cloud_normals_kernel<<<blocks_for(count, kThreads), kThreads, 0, (cudaStream_t) stream>>>(point_positions, image_height, image_width, normals);
	NNRT_LAUNCH_CHECK();
	return kOk;
}

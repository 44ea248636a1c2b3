// Shared device-side arithmetic for the sm_100a dense-depth GN kernels.
//
// Every index-deciding expression (anchor distances, NDC projection, edge functions, depth) is written
// in the reference's evaluation order and this translation unit is compiled with -fmad=false, so the
// rounding sequence equals the reference's x86-64 CPU build (no FMA contraction):
//   * 3-vector reductions are x + (y + z)  (Eigen redux order used by squaredNorm/dot/Matrix3f*Vector3f)
//   * Open3D TransformIndexer::Project is (fx*x)*(1/z) + cx
// References: cpp/rendering/functional/kernel/BarycentricCoordinates.h, cpp/rendering/kernel/RayFaceIntersection.h,
// cpp/rendering/kernel/CoordinateSystemConversions.h, cpp/core/kernel/KnnUtilities.h.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace nnrt_b200 {

constexpr float kEps = 1e-8f;           // K_EPSILON, cpp/rendering/kernel/RasterizationConstants.h:27
constexpr int kMaxAnchors = 8;          // MAX_ANCHOR_COUNT, cpp/geometry/functional/kernel/Defines.h:18
constexpr int kMaxFacesPerPixel = 8;    // MAX_POINTS_PER_PIXEL, cpp/rendering/kernel/RasterizationConstants.h:22
constexpr unsigned long long kEmptyKey = 0xFFFFFFFFFFFFFFFFull;

// status codes of the C ABI (include/nnrt_b200.h)
enum Status : int {
	kOk = 0, kInvalidArgument = 1, kCudaError = 2, kQueueOverflow = 3, kNotPositiveDefinite = 4, kUnsupported = 5
};

struct Intrinsics { float fx, fy, cx, cy; };
struct Rigid { float e[12]; };  // 3x4 row-major

__device__ __forceinline__ float sum3(float a, float b, float c) { return a + (b + c); }
__device__ __forceinline__ float dot3(float ax, float ay, float az, float bx, float by, float bz) {
	return sum3(ax * bx, ay * by, az * bz);
}
__device__ __forceinline__ float3 matvec3(const float* __restrict__ R, float3 v) {
	return make_float3(sum3(R[0] * v.x, R[1] * v.y, R[2] * v.z), sum3(R[3] * v.x, R[4] * v.y, R[5] * v.z),
	                   sum3(R[6] * v.x, R[7] * v.y, R[8] * v.z));
}
__device__ __forceinline__ float3 operator-(float3 a, float3 b) { return make_float3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ float3 operator+(float3 a, float3 b) { return make_float3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float3 operator*(float s, float3 a) { return make_float3(s * a.x, s * a.y, s * a.z); }

__device__ __forceinline__ float ndc_range(int dim1, int dim2) {
	float range = 2.0f;
	if (dim1 > dim2) range = __fdiv_rn(__fmul_rn((float) dim1, range), (float) dim2);
	return range;
}
// pixel centre -> NDC along one dimension (ImageSpacePixelAlongDimensionToNdc)
__device__ __forceinline__ float pixel_to_ndc(int i, int dim1, int dim2) {
	const float range = ndc_range(dim1, dim2);
	const float offset = __fmul_rn(range, 0.5f);  // == range / 2 exactly
	return __fadd_rn(-offset, __fdiv_rn(__fadd_rn(__fmul_rn(range, (float) i), offset), (float) dim1));
}
__device__ __forceinline__ float2 project(const Intrinsics& k, float3 p) {
	const float inv_z = __fdiv_rn(1.0f, p.z);
	return make_float2(__fadd_rn(__fmul_rn(__fmul_rn(k.fx, p.x), inv_z), k.cx), __fadd_rn(__fmul_rn(__fmul_rn(k.fy, p.y), inv_z), k.cy));
}
// the same with 1 / z already at hand (the fitter's warp kernel leaves it in the w component of every warped position: one division per
// vertex instead of one per face corner and per covered pixel; identical operations, identical bits)
__device__ __forceinline__ float2 project_with_inverse_depth(const Intrinsics& k, float x, float y, float inv_z) {
	return make_float2(__fadd_rn(__fmul_rn(__fmul_rn(k.fx, x), inv_z), k.cx), __fadd_rn(__fmul_rn(__fmul_rn(k.fy, y), inv_z), k.cy));
}
// Several numerators over ONE denominator, bit for bit the IEEE quotient (a zero quotient may differ in its sign only, -0 numerator
// over a positive divisor; nothing downstream tells the two zeros apart): r = RN(1 / d) once, then per numerator
//   q = RN(a r),  e = a - q d (exact in one FMA),  a / d = RN(q + e r)
// (Markstein's correction: with a correctly rounded reciprocal the corrected quotient is the correctly rounded one. It is the tail of the
// sequence the compiler emits for every `/` -- here the reciprocal and its refinement are shared, 3 instructions per quotient instead
// of ~10. A plain multiply by r is NOT enough: the half-ulp differences grow to 1e-3 .. 1e-2 in J^T J through the cancellations of the
// barycentric chain, DESIGN.md "Measured and dropped".) Valid for finite non-zero d and quotients in the normal range, which is what
// the Jacobian chain divides by (area + eps, area^2 + eps, depth, g^2 >= eps^2); checked against `/` on the device over 2^24 pairs
// by tests/test_gpu_stages.py::test_shared_divisor_quotient_is_the_ieee_quotient.
struct SharedDivisor { float d, r; };
__device__ __forceinline__ SharedDivisor shared_divisor(float d) { return SharedDivisor{d, __frcp_rn(d)}; }
__device__ __forceinline__ float operator/(float a, const SharedDivisor& s) {
	const float q = __fmul_rn(a, s.r);
	return __fmaf_rn(__fmaf_rn(-q, s.d, a), s.r, q);
}
__device__ __forceinline__ float3 rigid_transform(const Rigid& t, float3 p) {
	return make_float3(p.x * t.e[0] + p.y * t.e[1] + p.z * t.e[2] + t.e[3], p.x * t.e[4] + p.y * t.e[5] + p.z * t.e[6] + t.e[7],
	                   p.x * t.e[8] + p.y * t.e[9] + p.z * t.e[10] + t.e[11]);
}
__device__ __forceinline__ float3 rigid_rotate(const Rigid& t, float3 p) {
	return make_float3(p.x * t.e[0] + p.y * t.e[1] + p.z * t.e[2], p.x * t.e[4] + p.y * t.e[5] + p.z * t.e[6],
	                   p.x * t.e[8] + p.y * t.e[9] + p.z * t.e[10]);
}
// signed parallelogram area, clockwise-front convention
// (written with round-to-nearest intrinsics, like the rest of the per-pixel face test below: its results decide coverage and must
// not depend on the translation unit's -fmad / -prec-div settings)
__device__ __forceinline__ float edge_fn(float2 p, float2 v0, float2 v1) {
	return __fsub_rn(__fmul_rn(__fsub_rn(p.x, v0.x), __fsub_rn(v0.y, v1.y)), __fmul_rn(__fsub_rn(p.y, v0.y), __fsub_rn(v0.x, v1.x)));
}
__device__ __forceinline__ float fmin3(float a, float b, float c) { return fminf(fminf(a, b), c); }
__device__ __forceinline__ float fmax3(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }

__device__ __forceinline__ float point_segment_sq_distance(float2 p, float2 s0, float2 s1) {
	const float2 seg = make_float2(s1.x - s0.x, s1.y - s0.y);
	const float len2 = seg.x * seg.x + seg.y * seg.y;
	float ratio = (seg.x * (p.x - s0.x) + seg.y * (p.y - s0.y)) / len2;
	if (len2 <= kEps) {
		const float dx = p.x - s1.x, dy = p.y - s1.y;
		return dx * dx + dy * dy;
	}
	ratio = fminf(fmaxf(ratio, 0.f), 1.f);  // CPU saturatef of the reference (core/platform_independence/Algorithm.h)
	const float cx = s0.x + ratio * seg.x, cy = s0.y + ratio * seg.y;
	const float dx = cx - p.x, dy = cy - p.y;
	return dx * dx + dy * dy;
}

struct FaceNdc {
	float2 v0, v1, v2;
	float z0, z1, z2;
};

struct Fragment {
	float depth, distance;
	float3 bary;
};

// True when every point of the face's blur-expanded bounding box is closer (squared distance, as the reference compares it) than
// blur_radius to a vertex, hence to an edge: the reference's outside-pixel test `dist >= blur_radius` can then never reject a
// pixel of the box, and the three point-segment distances need not be evaluated to reproduce its decision. With the reference's
// blur radius (0.5 px) compared against a SQUARED distance this holds for every face smaller than about 10 pixels.
__device__ __forceinline__ bool face_box_within_blur(const FaceNdc& f, float blur_radius) {
	const float dx = (fmax3(f.v0.x, f.v1.x, f.v2.x) - fmin3(f.v0.x, f.v1.x, f.v2.x)) + 2.f * blur_radius;
	const float dy = (fmax3(f.v0.y, f.v1.y, f.v2.y) - fmin3(f.v0.y, f.v1.y, f.v2.y)) + 2.f * blur_radius;
	return (dx * dx + dy * dy) * 1.01f < blur_radius;  // 1 % margin over the fp32 rounding of both sides
}

// Per-pixel face test (UpdateQueueIfPixelInsideFace without the queue). Returns false if rejected. `within_blur`: the caller
// knows the distance test cannot reject this pixel (face_box_within_blur, or a pixel that already passed); out.distance is then 0.
__device__ __forceinline__ bool shade_pixel(const FaceNdc& f, float2 pixel, float blur_radius, bool perspective_correct,
                                            bool clip_bary, bool cull_back_faces, Fragment& out, bool within_blur = false) {
	const float face_area = edge_fn(f.v0, f.v1, f.v2);
	const bool is_back_face = face_area < 0.f;
	const bool zero_area = (face_area <= kEps && face_area >= -1.f * kEps);
	const float x_min = fmin3(f.v0.x, f.v1.x, f.v2.x) - blur_radius, x_max = fmax3(f.v0.x, f.v1.x, f.v2.x) + blur_radius;
	const float y_min = fmin3(f.v0.y, f.v1.y, f.v2.y) - blur_radius, y_max = fmax3(f.v0.y, f.v1.y, f.v2.y) + blur_radius;
	const bool z_invalid = fmax3(f.z0, f.z1, f.z2) < kEps;
	if (pixel.x > x_max || pixel.x < x_min || pixel.y > y_max || pixel.y < y_min || z_invalid) return false;
	if ((cull_back_faces && is_back_face) || zero_area) return false;
	// (both triples of quotients share a denominator: SharedDivisor above, the IEEE quotient's bits; the face's reciprocal is
	// loop-invariant in the callers' pixel loops)
	const SharedDivisor area_eps = shared_divisor(__fadd_rn(face_area, kEps));
	float3 bary = make_float3(edge_fn(pixel, f.v1, f.v2) / area_eps, edge_fn(pixel, f.v2, f.v0) / area_eps, edge_fn(pixel, f.v0, f.v1) / area_eps);
	if (perspective_correct) {
		const float n0 = __fmul_rn(__fmul_rn(bary.x, f.z1), f.z2);
		const float n1 = __fmul_rn(__fmul_rn(f.z0, bary.y), f.z2);
		const float n2 = __fmul_rn(__fmul_rn(f.z0, f.z1), bary.z);
		const SharedDivisor den = shared_divisor(fmaxf(__fadd_rn(__fadd_rn(n0, n1), n2), kEps));
		bary = make_float3(n0 / den, n1 / den, n2 / den);
	}
	float3 bc = bary;
	if (clip_bary) {
		bc = make_float3(fmaxf(bary.x, 0.f), fmaxf(bary.y, 0.f), fmaxf(bary.z, 0.f));
		const float n2 = sum3(bc.x * bc.x, bc.y * bc.y, bc.z * bc.z);
		if (n2 > 0.f) { const float n = sqrtf(n2); bc = make_float3(bc.x / n, bc.y / n, bc.z / n); }
	}
	const float depth = __fadd_rn(__fadd_rn(__fmul_rn(bc.x, f.z0), __fmul_rn(bc.y, f.z1)), __fmul_rn(bc.z, f.z2));
	if (depth < 0.f) return false;
	out.depth = depth;
	out.bary = bc;
	out.distance = 0.f;
	if (within_blur) return true;
	const float dist = fmin3(point_segment_sq_distance(pixel, f.v0, f.v1), point_segment_sq_distance(pixel, f.v0, f.v2),
	                         point_segment_sq_distance(pixel, f.v1, f.v2));
	const bool inside = bary.x > 0.f && bary.y > 0.f && bary.z > 0.f;
	if (!inside && dist >= blur_radius) return false;  // sic: squared distance against the radius
	out.distance = inside ? -dist : dist;
	return true;
}

// (depth, face) packed so that unsigned 64-bit order == the reference's RayFaceIntersection operator<
__device__ __forceinline__ unsigned long long pack_key(float depth, int face) {
	return ((unsigned long long) __float_as_uint(depth + 0.0f) << 32) | (unsigned int) face;
}

}  // namespace nnrt_b200

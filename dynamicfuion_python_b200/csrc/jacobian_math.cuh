// Per-pixel Jacobian chain of the dense-depth data term, kept entirely in registers.
//   d rho / d V  (3x9): barycentric coordinates w.r.t. the face's camera-space vertices
//   (cpp/alignment/functional/kernel/BarycentricCoordinateJacobians.h:43-419, ProjectionJacobians.h:26-39)
//   d w_l / d V = [v0 v1 v2] d rho/dV + rho (x) I3 ; d n_l / d V = [n0 n1 n2] d rho/dV
//   (cpp/alignment/functional/kernel/RasterizedSurfaceJacobiansImpl.h:43-201)
#pragma once
#include "common.cuh"

namespace nnrt_b200 {

struct BaryJacobian { float m[3][9]; };

// `stored_bary` is only read when !perspective_correct (sub-areas are then recovered from the stored coordinates).
// `p` are the projections of `v` (common.cuh::project) and `inverse_depth` the correctly rounded 1 / v[k].z they were made with: a caller
// that has already projected the face (the fused assembly: shade_pixel's face) hands them in, and the face area, the sub-areas and
// their quotients below become common subexpressions of its own.
__device__ __forceinline__ void barycentric_jacobian(BaryJacobian& J, float2 ray, const float3 v[3], const float2 p[3], const float inverse_depth[3],
                                                     const Intrinsics& ndc, bool perspective_correct, const float* stored_bary) {
	const float area = edge_fn(p[0], p[1], p[2]) + kEps;
	const float area_sq = area * area;
	// every denominator below is shared by several quotients: common.cuh::SharedDivisor (the IEEE quotient's bits, a third of its instructions)
	const SharedDivisor over_area = shared_divisor(area);
	float sub[3], distorted[3];
	if (perspective_correct) {
		sub[0] = edge_fn(ray, p[1], p[2]); sub[1] = edge_fn(ray, p[2], p[0]); sub[2] = edge_fn(ray, p[0], p[1]);
#pragma unroll
		for (int k = 0; k < 3; k++) distorted[k] = sub[k] / over_area;
	} else {
#pragma unroll
		for (int k = 0; k < 3; k++) { distorted[k] = stored_bary[k]; sub[k] = stored_bary[k] * area; }
	}
	// d(face area)/d p_i and d(sub area_r)/d p_i  (2-vectors)
	const float2 dA[3] = {make_float2(p[1].y - p[2].y, p[2].x - p[1].x), make_float2(p[2].y - p[0].y, p[0].x - p[2].x),
	                      make_float2(p[0].y - p[1].y, p[1].x - p[0].x)};
	// dS[r][i]: derivative of sub-area r w.r.t. ndc vertex i (zero for i == r)
	float2 dS[3][3];
#pragma unroll
	for (int r = 0; r < 3; r++) {
		const int a = (r + 1) % 3, b = (r + 2) % 3;  // sub-area r = edge_fn(ray, p[a], p[b])
		dS[r][r] = make_float2(0.f, 0.f);
		dS[r][a] = make_float2(p[b].y - ray.y, ray.x - p[b].x);
		dS[r][b] = make_float2(ray.y - p[a].y, p[a].x - ray.x);
	}
	const SharedDivisor den = shared_divisor(area_sq + kEps);
#pragma unroll
	for (int i = 0; i < 3; i++) {
		// 2x3 projection Jacobian of vertex i
		const SharedDivisor z = SharedDivisor{v[i].z, inverse_depth[i]}, z_sq = shared_divisor(v[i].z * v[i].z);
		const float px0 = ndc.fx / z, px2 = -ndc.fx * v[i].x / z_sq;
		const float py1 = ndc.fy / z, py2 = -ndc.fy * v[i].y / z_sq;
#pragma unroll
		for (int r = 0; r < 3; r++) {
			float dx, dy;
			if (i == r) {
				dx = (-sub[r] * dA[i].x) / den;
				dy = (-sub[r] * dA[i].y) / den;
			} else {
				dx = (area * dS[r][i].x - sub[r] * dA[i].x) / den;
				dy = (area * dS[r][i].y - sub[r] * dA[i].y) / den;
			}
			J.m[r][i * 3 + 0] = dx * px0 + dy * 0.f;
			J.m[r][i * 3 + 1] = dx * 0.f + dy * py1;
			J.m[r][i * 3 + 2] = dx * px2 + dy * py2;
		}
	}
	if (perspective_correct) {
		const float z0 = v[0].z, z1 = v[1].z, z2 = v[2].z;
		const float z12 = z1 * z2, z02 = z0 * z2, z01 = z0 * z1;
		const float n0 = distorted[0] * z12, n1 = distorted[1] * z02, n2 = distorted[2] * z01;
		const float g = fmaxf(n0 + n1 + n2, kEps);
		const SharedDivisor g_sq = shared_divisor(g * g);
		const float Dd[3][3] = {{(g - n0) * z12 / g_sq, -n0 * z02 / g_sq, -n0 * z01 / g_sq},
		                        {-n1 * z12 / g_sq, (g - n1) * z02 / g_sq, -n1 * z01 / g_sq},
		                        {-n2 * z12 / g_sq, -n2 * z02 / g_sq, (g - n2) * z01 / g_sq}};
		const float gz0 = distorted[1] * z2 + z1 * distorted[2];
		const float gz1 = distorted[0] * z2 + z0 * distorted[2];
		const float gz2 = distorted[0] * z1 + z0 * distorted[1];
		const float Dz[3][3] = {{-n0 * gz0 / g_sq, (g * distorted[0] * z2 - n0 * gz1) / g_sq, (g * distorted[0] * z1 - n0 * gz2) / g_sq},
		                        {(g * distorted[1] * z2 - n1 * gz0) / g_sq, -n1 * gz1 / g_sq, (g * distorted[1] * z0 - n1 * gz2) / g_sq},
		                        {(g * distorted[2] * z1 - n2 * gz0) / g_sq, (g * distorted[2] * z0 - n2 * gz1) / g_sq, -n2 * gz2 / g_sq}};
#pragma unroll
		for (int c = 0; c < 9; c++) {
			const float a0 = J.m[0][c], a1 = J.m[1][c], a2 = J.m[2][c];
#pragma unroll
			for (int r = 0; r < 3; r++) J.m[r][c] = sum3(Dd[r][0] * a0, Dd[r][1] * a1, Dd[r][2] * a2);
		}
#pragma unroll
		for (int r = 0; r < 3; r++)
#pragma unroll
			for (int i = 0; i < 3; i++) J.m[r][i * 3 + 2] += Dz[r][i];
	}
}

// the same from the camera-space vertices alone
__device__ __forceinline__ void barycentric_jacobian(BaryJacobian& J, float2 ray, const float3 v[3], const Intrinsics& ndc,
                                                     bool perspective_correct, const float* stored_bary) {
	float2 p[3];
	float inverse_depth[3];
#pragma unroll
	for (int k = 0; k < 3; k++) {
		inverse_depth[k] = __fdiv_rn(1.0f, v[k].z);
		p[k] = project_with_inverse_depth(ndc, v[k].x, v[k].y, inverse_depth[k]);
	}
	barycentric_jacobian(J, ray, v, p, inverse_depth, ndc, perspective_correct, stored_bary);
}

// out[r][c] = sum_k M_k[r] * J[k][c]  where M_k are the three column vectors (face vertices or normals)
__device__ __forceinline__ float col_mix(const float3 m[3], int r, const BaryJacobian& J, int c) {
	const float a0 = (r == 0 ? m[0].x : r == 1 ? m[0].y : m[0].z);
	const float a1 = (r == 0 ? m[1].x : r == 1 ? m[1].y : m[1].z);
	const float a2 = (r == 0 ? m[2].x : r == 1 ? m[2].y : m[2].z);
	return sum3(a0 * J.m[0][c], a1 * J.m[1][c], a2 * J.m[2][c]);
}

}  // namespace nnrt_b200

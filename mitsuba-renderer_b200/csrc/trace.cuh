// Ray/scene intersection device code. Replaces ShapeKDTree::rayIntersect (closest and shadow
// variants, src/librender/skdtree.cpp:108-132,180-199), SAHKDTree3D::rayIntersectHavran
// (include/mitsuba/render/sahkdtree3.h:170-300) and TriAccel::rayIntersect (triaccel.h:98-159).
//
// Parity design: WHICH triangles get tested differs from the kd-tree (we walk a 4-wide BVH), but
// every quantity that decides the result is computed with the reference's exact fp32 operation
// sequence — the scene-box clip, the adaptive epsilon, the Wald projection test, the inclusive
// [mint, maxt] interval and "a later equal t replaces an earlier one". Hence t/u/v are bit-equal
// and ids agree except when two triangles yield a bit-identical t (documented tie class).
#pragma once
#include "types.cuh"

#define TRACE_BLOCK        128
#define TRACE_SMEM_STACK   20      // entries per thread kept in shared memory
#define TRACE_LOCAL_STACK  76      // overflow entries in local memory (deep LBVH chains)

struct RayIn { float ox, oy, oz, mint, dx, dy, dz, maxt; };

// AABB::rayIntersect (include/mitsuba/core/aabb.h:349-382) on the enlarged kd-tree box, using
// dRcp = 1/d (ray.h:78-91), then the interval set-up of skdtree.cpp:114-124 / 184-196.
// Returns false when the reference would not traverse at all.
template <bool SHADOW>
__device__ __forceinline__ bool clipToScene(const DeviceScene &sc, const RayIn &r, float &mint, float &maxt) {
	float nearT = -INFINITY, farT = INFINITY;
	const float o[3] = { r.ox, r.oy, r.oz }, d[3] = { r.dx, r.dy, r.dz };
#pragma unroll
	for (int i = 0; i < 3; ++i) {
		const float minVal = sc.aabbMin[i], maxVal = sc.aabbMax[i];
		if (d[i] == 0) {
			if (o[i] < minVal || o[i] > maxVal) return false;
		} else {
			const float rcp = __fdiv_rn(1.0f, d[i]);
			float t1 = __fmul_rn(__fsub_rn(minVal, o[i]), rcp);
			float t2 = __fmul_rn(__fsub_rn(maxVal, o[i]), rcp);
			if (t1 > t2) { const float tmp = t1; t1 = t2; t2 = tmp; }
			nearT = (nearT < t1) ? t1 : nearT;        // std::max(nearT, t1)
			farT = (t2 < farT) ? t2 : farT;           // std::min(farT, t2)
			if (nearT > farT) return false;
		}
	}
	float rayMinT = r.mint;
	if (rayMinT == MTS_EPSILON) {
		float m = fmaxf(fmaxf(fabsf(r.ox), fabsf(r.oy)), fabsf(r.oz));
		if (!SHADOW) m = fmaxf(m, MTS_EPSILON);     // the shadow variants omit this clamp (skdtree.cpp:186-189)
		rayMinT = __fmul_rn(rayMinT, m);
	}
	mint = nearT; maxt = farT;
	if (rayMinT > mint) mint = rayMinT;
	if (r.maxt < maxt) maxt = r.maxt;
	return maxt > mint;
}

// TriAccel::rayIntersect (triaccel.h:98-159), operation for operation.
__device__ __forceinline__ bool triAccelTest(const float4 &q0, const float4 &q1, const float4 &q2, uint32_t k,
		const RayIn &r, float mint, float maxt, float &u, float &v, float &t) {
	float o_u, o_v, o_k, d_u, d_v, d_k;
	switch (k) {
		case 0: o_u = r.oy; o_v = r.oz; o_k = r.ox; d_u = r.dy; d_v = r.dz; d_k = r.dx; break;
		case 1: o_u = r.oz; o_v = r.ox; o_k = r.oy; d_u = r.dz; d_v = r.dx; d_k = r.dy; break;
		case 2: o_u = r.ox; o_v = r.oy; o_k = r.oz; d_u = r.dx; d_v = r.dy; d_k = r.dz; break;
		default: return false;
	}
	const float n_u = q0.y, n_v = q0.z, n_d = q0.w;
	const float recip = __fdiv_rn(1.0f, __fadd_rn(__fadd_rn(__fmul_rn(d_u, n_u), __fmul_rn(d_v, n_v)), d_k));
	t = __fmul_rn(__fsub_rn(__fsub_rn(__fsub_rn(n_d, __fmul_rn(o_u, n_u)), __fmul_rn(o_v, n_v)), o_k), recip);
	if (t < mint || t > maxt) return false;
	const float hu = __fsub_rn(__fadd_rn(o_u, __fmul_rn(t, d_u)), q1.x);
	const float hv = __fsub_rn(__fadd_rn(o_v, __fmul_rn(t, d_v)), q1.y);
	u = __fadd_rn(__fmul_rn(hv, q1.z), __fmul_rn(hu, q1.w));
	v = __fadd_rn(__fmul_rn(hu, q2.x), __fmul_rn(hv, q2.y));
	return u >= 0 && v >= 0 && __fadd_rn(u, v) <= 1.0f;
}

struct HitOut { float t, u, v; uint32_t tri; };   // tri = index into the sorted TriAccel array

#define SSTACK(i) sstack[(i) * TRACE_BLOCK + threadIdx.x]

__device__ __forceinline__ float safeRcp(float d) {
	// finite stand-in for 1/0 so that the slab test never produces NaN; the boxes are padded,
	// which makes the result conservative for axis-parallel rays
	return 1.0f / (fabsf(d) > 1e-20f ? d : copysignf(1e-20f, d));
}

// One ray against the 4-wide BVH. `sstack` is the block's shared stack (TRACE_SMEM_STACK *
// TRACE_BLOCK uint2, interleaved by thread so a warp's accesses are conflict-free); entries
// beyond it spill to local memory. Stack entries carry the child's entry distance so that
// nodes lying beyond the current closest hit are dropped when popped.
template <bool SHADOW, bool COUNT>
__device__ __forceinline__ bool traverse(const DeviceScene &sc, const RayIn &r, float mint, float maxt,
		uint2 *sstack, HitOut &hit, uint32_t &nodeVisits, uint32_t &triTests) {
	uint2 lstack[TRACE_LOCAL_STACK];
	int sp = 0;
	const float idx = safeRcp(r.dx), idy = safeRcp(r.dy), idz = safeRcp(r.dz);
	const float oix = r.ox * idx, oiy = r.oy * idy, oiz = r.oz * idz;
	bool found = false;
	uint32_t cur = 0;                 // root node
	hit.t = INFINITY; hit.u = hit.v = 0; hit.tri = 0xFFFFFFFFu;

	while (true) {
		if (!(cur & BVH_LEAF)) {
			if (COUNT) nodeVisits++;
			const float4 *np = sc.nodes + (size_t) cur * 8;
			const float4 lox = __ldg(np + 0), loy = __ldg(np + 1), loz = __ldg(np + 2);
			const float4 hix = __ldg(np + 3), hiy = __ldg(np + 4), hiz = __ldg(np + 5);
			const uint4 child = __ldg(reinterpret_cast<const uint4 *>(np + 6));
			float tn[4]; uint32_t cw[4] = { child.x, child.y, child.z, child.w };
			const float lx[4] = { lox.x, lox.y, lox.z, lox.w }, ly[4] = { loy.x, loy.y, loy.z, loy.w }, lz[4] = { loz.x, loz.y, loz.z, loz.w };
			const float hx[4] = { hix.x, hix.y, hix.z, hix.w }, hy[4] = { hiy.x, hiy.y, hiy.z, hiy.w }, hz[4] = { hiz.x, hiz.y, hiz.z, hiz.w };
#pragma unroll
			for (int i = 0; i < 4; ++i) {
				const float t0x = fmaf(lx[i], idx, -oix), t1x = fmaf(hx[i], idx, -oix);
				const float t0y = fmaf(ly[i], idy, -oiy), t1y = fmaf(hy[i], idy, -oiy);
				const float t0z = fmaf(lz[i], idz, -oiz), t1z = fmaf(hz[i], idz, -oiz);
				const float tnear = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fmaxf(fminf(t0z, t1z), mint));
				const float tfar = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fminf(fmaxf(t0z, t1z), maxt));
				tn[i] = (cw[i] != BVH_EMPTY && tnear <= tfar) ? tnear : INFINITY;
			}
			if (!SHADOW) {
				// sort the four (tn, cw) pairs ascending: 5-comparator network
#define CSWAP(a, b) if (tn[b] < tn[a]) { float tt = tn[a]; tn[a] = tn[b]; tn[b] = tt; uint32_t cc = cw[a]; cw[a] = cw[b]; cw[b] = cc; }
				CSWAP(0, 1) CSWAP(2, 3) CSWAP(0, 2) CSWAP(1, 3) CSWAP(1, 2)
#undef CSWAP
#pragma unroll
				for (int i = 3; i >= 1; --i) {
					if (tn[i] != INFINITY) {
						const uint2 e = make_uint2(cw[i], __float_as_uint(tn[i]));
						if (sp < TRACE_SMEM_STACK) SSTACK(sp) = e; else lstack[min(sp - TRACE_SMEM_STACK, TRACE_LOCAL_STACK - 1)] = e;
						sp++;
					}
				}
				if (tn[0] != INFINITY) { cur = cw[0]; continue; }
			} else {
				// any-hit: order is irrelevant; take the first hit child, push the rest
				uint32_t next = BVH_EMPTY;
#pragma unroll
				for (int i = 0; i < 4; ++i) {
					if (tn[i] != INFINITY) {
						if (next == BVH_EMPTY) next = cw[i];
						else {
							const uint2 e = make_uint2(cw[i], 0u);
							if (sp < TRACE_SMEM_STACK) SSTACK(sp) = e; else lstack[min(sp - TRACE_SMEM_STACK, TRACE_LOCAL_STACK - 1)] = e;
							sp++;
						}
					}
				}
				if (next != BVH_EMPTY) { cur = next; continue; }
			}
		} else {
			// leaf: (count-1)<<28 | first
			const uint32_t first = cur & 0x0FFFFFFFu, count = ((cur >> 28) & 7u) + 1u;
			for (uint32_t i = 0; i < count; ++i) {
				const float4 *tp = sc.triAccel + (size_t) (first + i) * 3;
				const float4 q0 = __ldg(tp), q1 = __ldg(tp + 1), q2 = __ldg(tp + 2);
				const uint32_t kraw = __float_as_uint(q0.x);
				if (SHADOW && (kraw & TRI_NONOCCLUDER)) continue;
				if (COUNT) triTests++;
				float u, v, t;
				if (triAccelTest(q0, q1, q2, kraw & TRI_K_MASK, r, mint, maxt, u, v, t)) {
					if (SHADOW) return true;
					maxt = t; found = true;            // sahkdtree3.h:278-283: later equal t replaces earlier
					hit.t = t; hit.u = u; hit.v = v; hit.tri = first + i;
				}
			}
		}
		// pop; drop entries that start strictly beyond the current closest hit
		while (true) {
			if (sp == 0) return found;
			--sp;
			const uint2 e = (sp < TRACE_SMEM_STACK) ? SSTACK(sp) : lstack[min(sp - TRACE_SMEM_STACK, TRACE_LOCAL_STACK - 1)];
			if (!SHADOW && __uint_as_float(e.y) > maxt) continue;
			cur = e.x;
			break;
		}
	}
}

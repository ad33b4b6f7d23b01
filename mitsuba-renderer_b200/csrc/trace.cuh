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
#define MTSGPU_HAVE_TRACE 1

#define TRACE_BLOCK        128
#ifndef TRACE_SMEM_STACK
#define TRACE_SMEM_STACK   20      // entries per thread kept in shared memory
#endif
#define TRACE_LOCAL_STACK  (96 - TRACE_SMEM_STACK)      // overflow entries in local memory (deep chains)

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
			const float rcp = __frcp_rn(d[i]);                 // correctly rounded, bit-identical to 1.0f / d
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
	const float recip = __frcp_rn(__fadd_rn(__fadd_rn(__fmul_rn(d_u, n_u), __fmul_rn(d_v, n_v)), d_k));   // == 1.0f / x, correctly rounded
	t = __fmul_rn(__fsub_rn(__fsub_rn(__fsub_rn(n_d, __fmul_rn(o_u, n_u)), __fmul_rn(o_v, n_v)), o_k), recip);
	if (t < mint || t > maxt) return false;
	const float hu = __fsub_rn(__fadd_rn(o_u, __fmul_rn(t, d_u)), q1.x);
	const float hv = __fsub_rn(__fadd_rn(o_v, __fmul_rn(t, d_v)), q1.y);
	u = __fadd_rn(__fmul_rn(hv, q1.z), __fmul_rn(hu, q1.w));
	v = __fadd_rn(__fmul_rn(hu, q2.x), __fmul_rn(hv, q2.y));
	return u >= 0 && v >= 0 && __fadd_rn(u, v) <= 1.0f;
}

// Analytic sphere stored in a TriAccel slot (k = 3, the low bits of the reference's KNoTriangleFlag, triaccel.h:27):
// {k, centre.xyz} {radius, ...}. Sphere::rayIntersect (src/shapes/sphere.cpp:92-131) with solveQuadratic
// (src/libcore/util.cpp:450-488), fp32 operation by operation; the closest-hit and the shadow version differ in
// how they treat a ray interval that starts inside the sphere.
#define TRI_K_SPHERE 3u
template <bool SHADOW>
__device__ __forceinline__ bool sphereTest(const float4 &q0, const float4 &q1, const RayIn &r, float mint, float maxt, float &t) {
	const float ox = __fsub_rn(r.ox, q0.y), oy = __fsub_rn(r.oy, q0.z), oz = __fsub_rn(r.oz, q0.w);
	const float radius = q1.x;
	const float A = __fadd_rn(__fadd_rn(__fmul_rn(r.dx, r.dx), __fmul_rn(r.dy, r.dy)), __fmul_rn(r.dz, r.dz));
	const float B = __fmul_rn(2.0f, __fadd_rn(__fadd_rn(__fmul_rn(r.dx, ox), __fmul_rn(r.dy, oy)), __fmul_rn(r.dz, oz)));
	const float C = __fsub_rn(__fadd_rn(__fadd_rn(__fmul_rn(ox, ox), __fmul_rn(oy, oy)), __fmul_rn(oz, oz)), __fmul_rn(radius, radius));
	float nearT, farT;
	if (A == 0) {
		if (B == 0) return false;
		nearT = farT = __fdiv_rn(-C, B);
	} else {
		const float discrim = __fsub_rn(__fmul_rn(B, B), __fmul_rn(__fmul_rn(4.0f, A), C));
		if (discrim < 0) return false;
		const float sqrtDiscrim = __fsqrt_rn(discrim);
		const float temp = (B < 0) ? __fmul_rn(-0.5f, __fsub_rn(B, sqrtDiscrim)) : __fmul_rn(-0.5f, __fadd_rn(B, sqrtDiscrim));
		nearT = __fdiv_rn(temp, A);
		farT = __fdiv_rn(C, temp);
		if (nearT > farT) { const float x = nearT; nearT = farT; farT = x; }
	}
	if (nearT > maxt || farT < mint) return false;
	if (SHADOW) return !(nearT < mint && farT > maxt);
	if (nearT < mint) {
		if (farT > maxt) return false;
		t = farT;
	} else {
		t = nearT;
	}
	return true;
}

// 256-bit read-only load (LDG.E.256 on sm_100): one L1 wavefront per lane where two 128-bit loads cost two.
// Written as v4.u64: ptxas 12.9 crashes on the v8.u32 form inside the persistent kernels.
__device__ __forceinline__ void ldg256(const void *p, uint32_t (&v)[8]) {
	unsigned long long a, b, c, d;
	asm("ld.global.nc.v4.u64 {%0, %1, %2, %3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p));
	v[0] = (uint32_t) a; v[1] = (uint32_t) (a >> 32); v[2] = (uint32_t) b; v[3] = (uint32_t) (b >> 32);
	v[4] = (uint32_t) c; v[5] = (uint32_t) (c >> 32); v[6] = (uint32_t) d; v[7] = (uint32_t) (d >> 32);
}

// Slab test of one ray against the four children of a quantised node (types.cuh: BVH4QNode): two 256-bit loads, then per
// plane one PRMT (byte -> the float 1 + q / 2^15) and one FFMA: t = m * S + C with S = scale / d and
// C = (origin - o) / d - S. The near and far words of each axis are picked by the ray's direction signs, so a child
// costs 6 PRMT + 6 FFMA + 4 min / max + 2 for the verdict. tn[i] = entry distance, +inf when child i is not entered
// (or empty). `one` = 0x3F800000 handed in from the kernel parameters (DeviceScene::oneBits): as a literal, ptxas made it
// PRMT's immediate and rebuilt the selector in a register before each of the 24 PRMTs (ncu source view, r02_q_c4_trace).
template <int I>
__device__ __forceinline__ float qPlane(uint32_t word, uint32_t one, float S, float C) {
	uint32_t m;
	asm("prmt.b32 %0, %1, %2, %3;" : "=r"(m) : "r"(word), "r"(one), "n"(0x7604 | (I << 4)));      // {0x00, byte I, 0x80, 0x3F}
	return fmaf(__uint_as_float(m), S, C);
}
// An empty slot needs no test of its own: the builder gives it an inverted box (never entered) and, should rounding ever
// collapse that box for a far-away degenerate node, the child word of slot 0 — a second visit of a real child changes no result.
__device__ __forceinline__ float entryOrInf(float tnear, float tfar) {
	float r;
	asm("{\n\t.reg .pred p;\n\tsetp.le.f32 p, %1, %2;\n\tselp.f32 %0, %1, 0f7F800000, p;\n\t}" : "=f"(r) : "f"(tnear), "f"(tfar));
	return r;
}
// The planes' t values carry the rounding of o / d (about 2^-22 |o / d| in all): for origins within a few scene extents that
// is far inside the boxes' pad, but not for a distant camera. Each axis is therefore widened by 2^-21 |o / d| (two FADDs
// per axis and node), so the box test stays conservative wherever the ray starts.
__device__ __forceinline__ void nodeTest(const float4 *nodes, uint32_t one, uint32_t node, float idx, float idy, float idz, float oix, float oiy, float oiz,
		float mint, float maxt, float (&tn)[4], uint32_t (&cw)[4]) {
	const char *np = reinterpret_cast<const char *>(nodes) + (size_t) node * 64;
	uint32_t a[8], b[8];
	ldg256(np, a); ldg256(np + 32, b);
	const float Sx = idx * __uint_as_float(a[3]), Sy = idy * __uint_as_float(b[6]), Sz = idz * __uint_as_float(b[7]);
	const float Cx = fmaf(__uint_as_float(a[0]), idx, -oix) - Sx, Cy = fmaf(__uint_as_float(a[1]), idy, -oiy) - Sy, Cz = fmaf(__uint_as_float(a[2]), idz, -oiz) - Sz;
	const float ex = fabsf(oix) * 4.76837158e-7f, ey = fabsf(oiy) * 4.76837158e-7f, ez = fabsf(oiz) * 4.76837158e-7f;      // 2^-21 |o / d|
	const float CxN = Cx - ex, CxF = Cx + ex, CyN = Cy - ey, CyF = Cy + ey, CzN = Cz - ez, CzF = Cz + ez;
	const bool nx = idx < 0.0f, ny = idy < 0.0f, nz = idz < 0.0f;
	const uint32_t nearX = nx ? b[3] : b[0], farX = nx ? b[0] : b[3];
	const uint32_t nearY = ny ? b[4] : b[1], farY = ny ? b[1] : b[4];
	const uint32_t nearZ = nz ? b[5] : b[2], farZ = nz ? b[2] : b[5];
#define NODE_CHILD(I) { \
		const float tnear = fmaxf(fmaxf(fmaxf(qPlane<I>(nearX, one, Sx, CxN), qPlane<I>(nearY, one, Sy, CyN)), qPlane<I>(nearZ, one, Sz, CzN)), mint); \
		const float tfar = fminf(fminf(fminf(qPlane<I>(farX, one, Sx, CxF), qPlane<I>(farY, one, Sy, CyF)), qPlane<I>(farZ, one, Sz, CzF)), maxt); \
		cw[I] = a[4 + I]; tn[I] = entryOrInf(tnear, tfar); }
	NODE_CHILD(0) NODE_CHILD(1) NODE_CHILD(2) NODE_CHILD(3)
#undef NODE_CHILD
}

// ---- instances of shape groups (src/shapes/instance.cpp) ---------------------------------------------------------
// A group's acceleration structure, walked with a stack in local memory only (the caller's shared stack is live).
// Same node layout, same interval rules and the same TriAccel test as traverse() below; groups hold triangles only.
#define GROUP_STACK 64
template <bool SHADOW>
__device__ __noinline__ bool traverseGroup(const DeviceScene &gs, const RayIn &r, float mint, float maxt, float &hitT, float &hitU, float &hitV, uint32_t &hitTri) {
	uint2 stack[GROUP_STACK];
	int sp = 0;
	const float dx = r.dx, dy = r.dy, dz = r.dz;
	const float idx = 1.0f / (fabsf(dx) > 1e-20f ? dx : copysignf(1e-20f, dx)), idy = 1.0f / (fabsf(dy) > 1e-20f ? dy : copysignf(1e-20f, dy)),
		idz = 1.0f / (fabsf(dz) > 1e-20f ? dz : copysignf(1e-20f, dz));
	const float oix = r.ox * idx, oiy = r.oy * idy, oiz = r.oz * idz;
	bool found = false;
	uint32_t cur = 0;                 // root node
	while (true) {
		if (!(cur & BVH_LEAF)) {
			float tn[4]; uint32_t cw[4];
			nodeTest(gs.nodes, gs.oneBits, cur, idx, idy, idz, oix, oiy, oiz, mint, maxt, tn, cw);
			uint32_t next = BVH_EMPTY; float nextT = INFINITY;
#pragma unroll
			for (int i = 0; i < 4; ++i) {
				if (tn[i] == INFINITY) continue;
				uint2 pushed;
				if (tn[i] < nextT) {                     // keep the nearest child, push the one it displaces
					if (next == BVH_EMPTY) { next = cw[i]; nextT = tn[i]; continue; }
					pushed = make_uint2(next, __float_as_uint(nextT));
					next = cw[i]; nextT = tn[i];
				} else pushed = make_uint2(cw[i], __float_as_uint(tn[i]));
				if (sp < GROUP_STACK && sp < gs.stackTotal) stack[sp++] = pushed; else atomicOr(gs.traceFlags, 1u);
			}
			if (next != BVH_EMPTY) { cur = next; continue; }
		} else {
			const uint32_t first = cur & 0x0FFFFFFFu, count = ((cur >> 28) & 7u) + 1u;
			for (uint32_t i = 0; i < count; ++i) {
				const float4 *tp = gs.triAccel + (size_t) (first + i) * 3;
				const float4 q0 = __ldg(tp), q1 = __ldg(tp + 1), q2 = __ldg(tp + 2);
				const uint32_t kraw = __float_as_uint(q0.x);
				if (SHADOW && (kraw & TRI_NONOCCLUDER)) continue;
				float u, v, t;
				if (triAccelTest(q0, q1, q2, kraw & TRI_K_MASK, r, mint, maxt, u, v, t)) {
					if (SHADOW) return true;
					maxt = t; found = true;
					hitT = t; hitU = u; hitV = v; hitTri = first + i;
				}
			}
		}
		while (true) {
			if (sp == 0) return found;
			const uint2 e = stack[--sp];
			if (!SHADOW && __uint_as_float(e.y) > maxt) continue;
			cur = e.x;
			break;
		}
	}
}

// Instance::rayIntersect (instance.cpp:72-86): the ray goes to object space through Transform::operator()(Ray)
// (transform.h:219-235: origin as a point, direction as a vector, NOT renormalised, so t keeps its meaning), then
// ShapeKDTree::rayIntersect(ray, mint, maxt, t, temp) (skdtree.h:435-466): clip to the group's box, tighten with the
// caller's interval, closest hit (or any hit) inside it.
template <bool SHADOW>
__device__ __noinline__ bool instanceTest(const InstanceDev *instances, const DeviceScene *groups, uint32_t instIndex, const RayIn r, float mint, float maxt,
		float &t, float &u, float &v, uint32_t &nested) {
	const InstanceDev &inst = instances[instIndex];
	const float *M = inst.w2o;
	RayIn o;
	o.ox = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(M[0], r.ox), __fmul_rn(M[1], r.oy)), __fmul_rn(M[2], r.oz)), M[3]);
	o.oy = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(M[4], r.ox), __fmul_rn(M[5], r.oy)), __fmul_rn(M[6], r.oz)), M[7]);
	o.oz = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(M[8], r.ox), __fmul_rn(M[9], r.oy)), __fmul_rn(M[10], r.oz)), M[11]);
	const float w = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(M[12], r.ox), __fmul_rn(M[13], r.oy)), __fmul_rn(M[14], r.oz)), M[15]);
	if (w != 1.0f) { const float rw = __frcp_rn(w); o.ox = __fmul_rn(o.ox, rw); o.oy = __fmul_rn(o.oy, rw); o.oz = __fmul_rn(o.oz, rw); }
	o.dx = __fadd_rn(__fadd_rn(__fmul_rn(M[0], r.dx), __fmul_rn(M[1], r.dy)), __fmul_rn(M[2], r.dz));
	o.dy = __fadd_rn(__fadd_rn(__fmul_rn(M[4], r.dx), __fmul_rn(M[5], r.dy)), __fmul_rn(M[6], r.dz));
	o.dz = __fadd_rn(__fadd_rn(__fmul_rn(M[8], r.dx), __fmul_rn(M[9], r.dy)), __fmul_rn(M[10], r.dz));
	o.mint = r.mint; o.maxt = r.maxt;
	const DeviceScene &gs = groups[inst.group];
	float nearT = -INFINITY, farT = INFINITY;
	const float oo[3] = { o.ox, o.oy, o.oz }, dd[3] = { o.dx, o.dy, o.dz };
#pragma unroll
	for (int i = 0; i < 3; ++i) {                              // AABB::rayIntersect (aabb.h:349-382)
		const float minVal = gs.aabbMin[i], maxVal = gs.aabbMax[i];
		if (dd[i] == 0) {
			if (oo[i] < minVal || oo[i] > maxVal) return false;
		} else {
			const float rcp = __frcp_rn(dd[i]);
			float t1 = __fmul_rn(__fsub_rn(minVal, oo[i]), rcp), t2 = __fmul_rn(__fsub_rn(maxVal, oo[i]), rcp);
			if (t1 > t2) { const float tmp = t1; t1 = t2; t2 = tmp; }
			nearT = (nearT < t1) ? t1 : nearT;
			farT = (t2 < farT) ? t2 : farT;
			if (nearT > farT) return false;
		}
	}
	if (mint > nearT) nearT = mint;
	if (maxt < farT) farT = maxt;
	if (!(farT > nearT)) return false;
	return traverseGroup<SHADOW>(gs, o, nearT, farT, t, u, v, nested);
}

// Cylinder::rayIntersect (src/shapes/cylinder.cpp:76-142): the ray in object space (Transform::operator()(Ray)), the
// quadratic of the infinite tube in x / y (solveQuadratic, util.cpp:450-488), then the z range of the two roots.
template <bool SHADOW>
__device__ __noinline__ bool cylinderTest(const InstanceDev *instances, uint32_t index, const RayIn r, float mint, float maxt, float &t) {
	const InstanceDev &c = instances[index];
	const float *M = c.w2o;
	float ox = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(M[0], r.ox), __fmul_rn(M[1], r.oy)), __fmul_rn(M[2], r.oz)), M[3]);
	float oy = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(M[4], r.ox), __fmul_rn(M[5], r.oy)), __fmul_rn(M[6], r.oz)), M[7]);
	float oz = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(M[8], r.ox), __fmul_rn(M[9], r.oy)), __fmul_rn(M[10], r.oz)), M[11]);
	const float w = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(M[12], r.ox), __fmul_rn(M[13], r.oy)), __fmul_rn(M[14], r.oz)), M[15]);
	if (w != 1.0f) { const float rw = __frcp_rn(w); ox = __fmul_rn(ox, rw); oy = __fmul_rn(oy, rw); oz = __fmul_rn(oz, rw); }
	const float dx = __fadd_rn(__fadd_rn(__fmul_rn(M[0], r.dx), __fmul_rn(M[1], r.dy)), __fmul_rn(M[2], r.dz));
	const float dy = __fadd_rn(__fadd_rn(__fmul_rn(M[4], r.dx), __fmul_rn(M[5], r.dy)), __fmul_rn(M[6], r.dz));
	const float dz = __fadd_rn(__fadd_rn(__fmul_rn(M[8], r.dx), __fmul_rn(M[9], r.dy)), __fmul_rn(M[10], r.dz));
	const float A = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
	const float B = __fmul_rn(2.0f, __fadd_rn(__fmul_rn(dx, ox), __fmul_rn(dy, oy)));
	const float C = __fsub_rn(__fadd_rn(__fmul_rn(ox, ox), __fmul_rn(oy, oy)), __fmul_rn(c.radius, c.radius));
	float nearT, farT;
	if (A == 0) {
		if (B == 0) return false;
		nearT = farT = __fdiv_rn(-C, B);
	} else {
		const float discrim = __fsub_rn(__fmul_rn(B, B), __fmul_rn(__fmul_rn(4.0f, A), C));
		if (discrim < 0) return false;
		const float sqrtDiscrim = __fsqrt_rn(discrim);
		const float temp = (B < 0) ? __fmul_rn(-0.5f, __fsub_rn(B, sqrtDiscrim)) : __fmul_rn(-0.5f, __fadd_rn(B, sqrtDiscrim));
		nearT = __fdiv_rn(temp, A);
		farT = __fdiv_rn(C, temp);
		if (nearT > farT) { const float x = nearT; nearT = farT; farT = x; }
	}
	if (nearT > maxt || farT < mint) return false;
	const float zPosNear = __fadd_rn(oz, __fmul_rn(dz, nearT));
	const float zPosFar = __fadd_rn(oz, __fmul_rn(dz, farT));
	if (zPosNear >= 0 && zPosNear <= c.length && nearT >= mint) {
		t = nearT;
		return true;
	} else if (zPosFar >= 0 && zPosFar <= c.length) {
		if (farT > maxt) return false;
		t = farT;
		return true;
	}
	return false;
}

template <bool SHADOW, bool SPH>
__device__ __forceinline__ bool primTest(const DeviceScene &sc, const float4 &q0, const float4 &q1, const float4 &q2, uint32_t kraw, const RayIn &r,
		float mint, float maxt, float &u, float &v, float &t, uint32_t &nested) {
	if (SPH && (kraw & TRI_K_MASK) == TRI_K_SPHERE) {
		u = v = 0.0f;
		if (q1.y == 2.0f) return cylinderTest<SHADOW>(sc.instances, __float_as_uint(q1.z), r, mint, maxt, t);          // a_v = 2: a cylinder
		if (q1.y != 0.0f) return instanceTest<SHADOW>(sc.instances, sc.groups, __float_as_uint(q1.z), r, mint, maxt, t, u, v, nested);   // a_v = 1: an instance
		return sphereTest<SHADOW>(q0, q1, r, mint, maxt, t);
	}
	return triAccelTest(q0, q1, q2, kraw & TRI_K_MASK, r, mint, maxt, u, v, t);
}

struct HitOut { float t, u, v; uint32_t tri; uint32_t nested; };   // tri = index into the sorted TriAccel array; nested: for an instance
                                                                  // hit, the sorted triangle index inside its group (else 0xFFFFFFFF)

#define SSTACK(i) sstack[(i) * TRACE_BLOCK + threadIdx.x]

// The same slots through a 32-bit shared-window address computed once per thread: the generic-pointer form above makes
// the compiler rebuild the window base (S2R SR_CgaCtaId + LEA) next to every predicated store of the push sequence.
__device__ __forceinline__ void stackStore(uint32_t base, int i, uint2 e) {
	asm volatile("st.shared.v2.u32 [%0], {%1, %2};" :: "r"(base + (uint32_t) i * (TRACE_BLOCK * 8u)), "r"(e.x), "r"(e.y));
}
__device__ __forceinline__ uint2 stackLoad(uint32_t base, int i) {
	uint2 e;
	asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(e.x), "=r"(e.y) : "r"(base + (uint32_t) i * (TRACE_BLOCK * 8u)));
	return e;
}

__device__ __forceinline__ float safeRcp(float d) {
	// finite stand-in for 1/0 so that the slab test never produces NaN; the boxes are padded,
	// which makes the result conservative for axis-parallel rays
	return 1.0f / (fabsf(d) > 1e-20f ? d : copysignf(1e-20f, d));
}

// One ray against the 4-wide BVH. `sstack` is the block's shared stack (TRACE_SMEM_STACK *
// TRACE_BLOCK uint2, interleaved by thread so a warp's accesses are conflict-free); entries
// beyond it spill to local memory. Stack entries carry the child's entry distance so that
// nodes lying beyond the current closest hit are dropped when popped.
template <bool SHADOW, bool COUNT, bool SPH = false>
__device__ __forceinline__ bool traverse(const DeviceScene &sc, const RayIn &r, float mint, float maxt,
		uint2 *sstack, HitOut &hit, uint32_t &nodeVisits, uint32_t &triTests) {
	uint2 lstack[TRACE_LOCAL_STACK];
	const uint32_t sbase = (uint32_t) __cvta_generic_to_shared(sstack) + threadIdx.x * 8u;
	int sp = 0;
	uint32_t deepest = 0;
	const float idx = safeRcp(r.dx), idy = safeRcp(r.dy), idz = safeRcp(r.dz);
	const float oix = r.ox * idx, oiy = r.oy * idy, oiz = r.oz * idz;
	bool found = false;
	uint32_t cur = 0;                 // root node
	hit.t = INFINITY; hit.u = hit.v = 0; hit.tri = 0xFFFFFFFFu; hit.nested = 0xFFFFFFFFu;

	while (true) {
		if (!(cur & BVH_LEAF)) {
			if (COUNT) nodeVisits++;
			float tn[4]; uint32_t cw[4];
			nodeTest(sc.nodes, sc.oneBits, cur, idx, idy, idz, oix, oiy, oiz, mint, maxt, tn, cw);
			{
				// nearest hit first; the others in any order (popped entries are culled by their entry distance)
#define CSWAP(a, b) if (tn[b] < tn[a]) { float tt = tn[a]; tn[a] = tn[b]; tn[b] = tt; uint32_t cc = cw[a]; cw[a] = cw[b]; cw[b] = cc; }
				CSWAP(0, 1) CSWAP(0, 2) CSWAP(0, 3)
#undef CSWAP
				const int h1 = tn[1] != INFINITY, h2 = tn[2] != INFINITY, h3 = tn[3] != INFINITY;
				if (sp + 3 <= sc.stackFast) {       // common case: three predicated stores, no branches
					if (h1) stackStore(sbase, sp, make_uint2(cw[1], __float_as_uint(tn[1])));
					if (h2) stackStore(sbase, sp + h1, make_uint2(cw[2], __float_as_uint(tn[2])));
					if (h3) stackStore(sbase, sp + h1 + h2, make_uint2(cw[3], __float_as_uint(tn[3])));
					sp += h1 + h2 + h3;
				} else {
#pragma unroll
					for (int i = 1; i <= 3; ++i) {
						if (tn[i] != INFINITY) {
							const uint2 e = make_uint2(cw[i], __float_as_uint(tn[i]));
							if (sp >= sc.stackTotal) { atomicOr(sc.traceFlags, 1u); continue; }       // stack full: reported by the host call (MTSGPU_ERR_LIMIT)
							if (sp < TRACE_SMEM_STACK) stackStore(sbase, sp, e); else lstack[sp - TRACE_SMEM_STACK] = e;
							sp++;
						}
					}
				}
				if (COUNT && (uint32_t) sp > deepest) { deepest = (uint32_t) sp; atomicMax(sc.traceFlags + 1, deepest); }      // counter builds: deepest stack seen
				if (tn[0] != INFINITY) { cur = cw[0]; continue; }
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
				uint32_t nested = 0xFFFFFFFFu;
				if (primTest<SHADOW, SPH>(sc, q0, q1, q2, kraw, r, mint, maxt, u, v, t, nested)) {
					if (SHADOW) return true;
					maxt = t; found = true;            // sahkdtree3.h:278-283: later equal t replaces earlier
					hit.t = t; hit.u = u; hit.v = v; hit.tri = first + i;
					if (SPH) hit.nested = nested;
				}
			}
		}
		// pop; drop entries that start strictly beyond the current closest hit
		while (true) {
			if (sp == 0) return found;
			--sp;
			const uint2 e = (sp < TRACE_SMEM_STACK) ? stackLoad(sbase, sp) : lstack[sp - TRACE_SMEM_STACK];
			if (!SHADOW && __uint_as_float(e.y) > maxt) continue;
			cur = e.x;
			break;
		}
	}
}

// =====================================================================================================
// Persistent-warp traversal with per-lane ray refetch.
//
// ncu on the one-thread-per-ray kernel (profiles/r01_trace_closest_bvh4_details.csv) showed 4.4 of 32
// lanes active per issued instruction on incoherent rays: lanes whose ray missed the scene box or finished
// early idle until the slowest ray of the warp is done. Here a fixed grid of warps pulls rays from a global
// counter (in chunks of TRACE_CHUNK per warp, handed out lane by lane with ballot + popcount); a lane that
// finishes its ray gets the next one at the top of the loop, so a warp stays full until the queue is empty.
// Each loop iteration performs one traversal step (one 4-wide node, or one leaf) per active lane.
// =====================================================================================================
#define TRACE_CHUNK 128
#define TRACE_PERSISTENT_BLOCKS_PER_SM 8
#define TRACE_FETCH_ATTEMPTS 8
#define TRACE_REFETCH_BELOW 28      // leave the step loop to refetch when fewer lanes than this are active
#define TRACE_LEAF_THRESHOLD 6      // lanes holding a leaf that end the inner-node phase

struct LaneState {
	RayIn r;
	float mint, maxt, idx, idy, idz, oix, oiy, oiz;
	uint32_t cur, leaf, ray, tri;     // cur: next node/leaf to visit (BVH_EMPTY: none); leaf: postponed leaf
	float t, u, v;
	int sp;
	bool active, found;
};

// Sink concept:  void closest(uint32_t ray, bool found, float t, float u, float v, uint32_t tri);
//                void any(uint32_t ray, bool occluded);
//
// Step structure ("while-while" with one postponed leaf, after Aila & Laine 2009): all lanes walk inner nodes
// until every lane holds a leaf (lanes that already hold one keep walking speculatively), then all lanes test
// their leaf's triangles together. The first version tested a leaf as soon as a lane reached it; ncu's source
// view showed the TriAccel test — 55 % of all issued instructions — running with 2 of 32 lanes active.
// The visiting order never changes the closest hit (only which of several bit-identical t values is kept).
template <bool SHADOW, bool COUNT, class Sink, bool SPH = false>
__device__ __forceinline__ void tracePersistent(const DeviceScene &sc, const float4 *__restrict__ rays, uint32_t n,
		uint32_t *counter, Sink &sink, uint2 *sstack, uint32_t &nodeVisits, uint32_t &triTests, int refetchBelow, bool speculate = true, int leafThreshold = 16) {
	uint2 lstack[TRACE_LOCAL_STACK];
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t ltMask = (1u << lane) - 1u;
	uint32_t poolNext = 0, poolEnd = 0;            // warp-uniform
	bool exhausted = false;                        // this lane drew an index >= n
	const uint32_t sbase = (uint32_t) __cvta_generic_to_shared(sstack) + threadIdx.x * 8u;
	uint32_t deepest = 0;
	LaneState s;
	s.active = false; s.sp = 0; s.cur = BVH_EMPTY; s.leaf = BVH_EMPTY; s.found = false;

#define STACK_PUSH(e) do { if (s.sp >= sc.stackTotal) atomicOr(sc.traceFlags, 1u);          /* stack full: reported by the host call (MTSGPU_ERR_LIMIT) */ \
		else { if (s.sp < TRACE_SMEM_STACK) stackStore(sbase, s.sp, (e)); else lstack[s.sp - TRACE_SMEM_STACK] = (e); s.sp++; } } while (0)
	// pop the next entry that can still matter; BVH_EMPTY when the stack is empty
#define STACK_POP(dst) do { (dst) = BVH_EMPTY; while (s.sp > 0) { --s.sp; \
		const uint2 e_ = (s.sp < TRACE_SMEM_STACK) ? stackLoad(sbase, s.sp) : lstack[s.sp - TRACE_SMEM_STACK]; \
		if (!SHADOW && __uint_as_float(e_.y) > s.maxt) continue; (dst) = e_.x; break; } } while (0)

	while (true) {
		// ---- refetch: lanes without a ray draw the next ray index; a ray that misses the scene box is
		//      answered on the spot, so try a few times before stepping
#pragma unroll 1
		for (int attempt = 0; attempt < TRACE_FETCH_ATTEMPTS; ++attempt) {
			const bool mine = !s.active && !exhausted;
			const uint32_t need = __ballot_sync(0xFFFFFFFFu, mine);
			if (need == 0) break;
			const uint32_t cnt = __popc(need), rank = __popc(need & ltMask);
			const uint32_t avail = poolEnd - poolNext;
			uint32_t idx = 0xFFFFFFFFu;
			if (mine && rank < avail) idx = poolNext + rank;
			if (cnt > avail) {
				uint32_t nb = 0;
				if (lane == 0) nb = atomicAdd(counter, (uint32_t) TRACE_CHUNK);
				nb = __shfl_sync(0xFFFFFFFFu, nb, 0);
				if (mine && rank >= avail) idx = nb + (rank - avail);
				poolNext = nb + (cnt - avail); poolEnd = nb + TRACE_CHUNK;
			} else {
				poolNext += cnt;
			}
			if (mine) {
				if (idx >= n) {
					exhausted = true;
				} else {
					const float4 a = __ldg(rays + 2 * (size_t) idx), b = __ldg(rays + 2 * (size_t) idx + 1);
					s.r.ox = a.x; s.r.oy = a.y; s.r.oz = a.z; s.r.mint = a.w; s.r.dx = b.x; s.r.dy = b.y; s.r.dz = b.z; s.r.maxt = b.w;
					s.ray = idx;
					if (clipToScene<SHADOW>(sc, s.r, s.mint, s.maxt)) {
						s.idx = safeRcp(s.r.dx); s.idy = safeRcp(s.r.dy); s.idz = safeRcp(s.r.dz);
						s.oix = s.r.ox * s.idx; s.oiy = s.r.oy * s.idy; s.oiz = s.r.oz * s.idz;
						s.cur = 0; s.leaf = BVH_EMPTY; s.sp = 0; s.found = false; s.active = true;
						s.t = INFINITY; s.u = s.v = 0; s.tri = 0xFFFFFFFFu;
					} else {
						if (SHADOW) sink.any(idx, false); else sink.closest(idx, false, INFINITY, 0.0f, 0.0f, 0xFFFFFFFFu);
					}
				}
			}
		}
		if (!__any_sync(0xFFFFFFFFu, s.active)) {
			if (__all_sync(0xFFFFFFFFu, exhausted)) break;
			continue;
		}
		// (A per-warp switch to late thresholds, 12 / 16, for nearly parallel rays paid off while every lane tested its own
		// leaf; with the shared leaf phase below coherent camera rays also run fastest at 28 / 6 — C4 camera rays 2.66 ->
		// 3.04 Grays/s, the 1M-triangle C2 camera batch 1.65 -> 1.77 ms — so one pair of thresholds serves all warps.)
		const int rbNow = refetchBelow, ltNow = leafThreshold;

		// ---- traversal; stay here while the warp is reasonably full (or no rays are left to fetch)
		while (true) {
			// (A) inner nodes, until every lane holds a leaf or has run out of nodes
			while (true) {
				const bool inner = s.active && !(s.cur & BVH_LEAF) && (speculate || s.leaf == BVH_EMPTY);      // BVH_EMPTY has the leaf bit set
				if (inner) {
					if (COUNT) nodeVisits++;
					float tn[4]; uint32_t cw[4];
					nodeTest(sc.nodes, sc.oneBits, s.cur, s.idx, s.idy, s.idz, s.oix, s.oiy, s.oiz, s.mint, s.maxt, tn, cw);
					{
						// bring the nearest hit to slot 0; the order of the others is immaterial (measured: 28.3 vs
						// 28.5 node visits per ray with a full sort) because popped entries are culled by distance
#define CSWAP(a, b) if (tn[b] < tn[a]) { float tt = tn[a]; tn[a] = tn[b]; tn[b] = tt; uint32_t cc = cw[a]; cw[a] = cw[b]; cw[b] = cc; }
						CSWAP(0, 1) CSWAP(0, 2) CSWAP(0, 3)
#undef CSWAP
					}
					{
						// push the remaining hits. Common case: room in the shared-memory stack -> three predicated
						// stores at computed slots, no branches
						const int h1 = tn[1] != INFINITY, h2 = tn[2] != INFINITY, h3 = tn[3] != INFINITY;
						if (s.sp + 3 <= sc.stackFast) {
							if (h1) stackStore(sbase, s.sp, make_uint2(cw[1], __float_as_uint(tn[1])));
							if (h2) stackStore(sbase, s.sp + h1, make_uint2(cw[2], __float_as_uint(tn[2])));
							if (h3) stackStore(sbase, s.sp + h1 + h2, make_uint2(cw[3], __float_as_uint(tn[3])));
							s.sp += h1 + h2 + h3;
						} else {
#pragma unroll
							for (int i = 1; i <= 3; ++i)
								if (tn[i] != INFINITY) STACK_PUSH(make_uint2(cw[i], __float_as_uint(tn[i])));
						}
					}
					if (COUNT && (uint32_t) s.sp > deepest) { deepest = (uint32_t) s.sp; atomicMax(sc.traceFlags + 1, deepest); }      // counter builds: deepest stack seen
					// next node: the nearest child, else the stack. A leaf is parked (one slot) and the walk goes on from the
					// stack. ONE pop site serves both cases: as two divergent regions the pops were 8.9 % of the issue slots at
					// ~3 active lanes (profiles/r02_last_c4_64spp_digest.txt); lanes that pop for either reason now share the issue.
					bool pop = tn[0] == INFINITY;
					if (!pop) {
						s.cur = cw[0];
						if ((s.cur & BVH_LEAF) && s.leaf == BVH_EMPTY) { s.leaf = s.cur; pop = true; }
					}
					while (pop) {
						STACK_POP(s.cur);
						pop = (s.cur & BVH_LEAF) && s.cur != BVH_EMPTY && s.leaf == BVH_EMPTY;
						if (pop) s.leaf = s.cur;
					}
				}
				const bool canWalk = s.active && !(s.cur & BVH_LEAF);
				if (!__any_sync(0xFFFFFFFFu, canWalk && s.leaf == BVH_EMPTY)) break;   // nobody without a leaf can go on
				if (__popc(__ballot_sync(0xFFFFFFFFu, s.active && s.leaf != BVH_EMPTY)) >= ltNow) break;   // enough leaves to test together
			}
			// (B) leaves. ncu's source view of the per-lane loop (profiles/r02_q2_*): 37 % of all issue slots at 3 to 5 active
			// lanes. Triangle variant (!SPH): the warp shares the work out. Up to 8 lanes that hold a leaf ("owners") post their
			// ray and leaf word in shared memory; lane w then tests triangle (w & 3) of owner (w >> 2); the four lanes of an owner
			// fold their results by shuffles — smallest t, of equal t the later triangle — which is the very result of the
			// sequential loop (sahkdtree3.h:278-283: a later equal t replaces an earlier one).
			if (!SPH) {
				static_assert(BVH_LEAF_MAX <= 4, "the shared leaf phase gives every owner four test slots");
				__shared__ float4 xRayA[TRACE_BLOCK / 32][8], xRayB[TRACE_BLOCK / 32][8];
				__shared__ uint32_t xLeaf[TRACE_BLOCK / 32][8];
				const uint32_t wid = threadIdx.x >> 5;
				const bool has = s.active && s.leaf != BVH_EMPTY;
				const uint32_t hasMask = __ballot_sync(0xFFFFFFFFu, has);
				const int rank = __popc(hasMask & ltMask), nOwners = __popc(hasMask);
				for (int base = 0; base < nOwners; base += 8) {
					const bool mine = has && rank >= base && rank < base + 8;
					if (mine) {
						xRayA[wid][rank - base] = make_float4(s.r.ox, s.r.oy, s.r.oz, s.mint);
						xRayB[wid][rank - base] = make_float4(s.r.dx, s.r.dy, s.r.dz, s.maxt);
						xLeaf[wid][rank - base] = s.leaf;
					}
					__syncwarp();
					const int slot = lane >> 2, k = lane & 3;
					float rt = 0.0f, ru = 0.0f, rv = 0.0f; int rk = -1;          // this lane's result; rk = triangle slot, -1: not accepted
					if (base + slot < nOwners) {
						const uint32_t leaf = xLeaf[wid][slot];
						if ((uint32_t) k <= ((leaf >> 28) & 7u)) {
							const float4 ra = xRayA[wid][slot], rb = xRayB[wid][slot];
							const float4 *tp = sc.triAccel + (size_t) ((leaf & 0x0FFFFFFFu) + k) * 3;
							const float4 q0 = __ldg(tp), q1 = __ldg(tp + 1), q2 = __ldg(tp + 2);
							const uint32_t kraw = __float_as_uint(q0.x);
							if (!(SHADOW && (kraw & TRI_NONOCCLUDER))) {
								if (COUNT) triTests++;
								RayIn rr; rr.ox = ra.x; rr.oy = ra.y; rr.oz = ra.z; rr.mint = 0; rr.dx = rb.x; rr.dy = rb.y; rr.dz = rb.z; rr.maxt = 0;
								if (triAccelTest(q0, q1, q2, kraw & TRI_K_MASK, rr, ra.w, rb.w, ru, rv, rt)) rk = k;
							}
						}
					}
					if (SHADOW) {
						// any-hit: an owner only needs to know whether one of its four lanes accepted
						const uint32_t acc = __ballot_sync(0xFFFFFFFFu, rk >= 0);
						if (mine && ((acc >> (4 * (rank - base))) & 0xFu)) s.found = true;
					} else {
						// closest hit: the four lanes of an owner fold their results with two butterfly steps — smallest t wins, of
						// equal t the LATER triangle (what the sequential loop's "t <= maxt, then maxt = t" leaves behind) — and the
						// owner fetches the winner with four shuffles
#pragma unroll
						for (int m = 1; m <= 2; m <<= 1) {
							const float ot = __shfl_xor_sync(0xFFFFFFFFu, rt, m), ou = __shfl_xor_sync(0xFFFFFFFFu, ru, m), ov = __shfl_xor_sync(0xFFFFFFFFu, rv, m);
							const int ok = __shfl_xor_sync(0xFFFFFFFFu, rk, m);
							if (ok >= 0 && (rk < 0 || ot < rt || (ot == rt && ok > rk))) { rt = ot; ru = ou; rv = ov; rk = ok; }
						}
						const int src = mine ? 4 * (rank - base) : 0;
						const float wt = __shfl_sync(0xFFFFFFFFu, rt, src), wu = __shfl_sync(0xFFFFFFFFu, ru, src), wv = __shfl_sync(0xFFFFFFFFu, rv, src);
						const int wk = __shfl_sync(0xFFFFFFFFu, rk, src);
						if (mine && wk >= 0) {          // accepted against the owner's [mint, maxt] as posted, i.e. wt <= s.maxt
							s.found = true;
							s.maxt = wt; s.t = wt; s.u = wu; s.v = wv; s.tri = (s.leaf & 0x0FFFFFFFu) + (uint32_t) wk;
						}
					}
					if (mine) {
						s.leaf = BVH_EMPTY;
						if (SHADOW && s.found) { s.sp = 0; s.cur = BVH_EMPTY; }          // any-hit: done
						else if ((s.cur & BVH_LEAF) && s.cur != BVH_EMPTY) { s.leaf = s.cur; STACK_POP(s.cur); }
					}
					__syncwarp();
				}
			} else
			// scenes with analytic shapes: every lane that holds a leaf tests its primitives; a second leaf waiting in `cur` moves up
			while (true) {
				const bool has = s.active && s.leaf != BVH_EMPTY;
				if (!__any_sync(0xFFFFFFFFu, has)) break;
				if (has) {
					const uint32_t first = s.leaf & 0x0FFFFFFFu, count = ((s.leaf >> 28) & 7u) + 1u;
					s.leaf = BVH_EMPTY;
					for (uint32_t i = 0; i < count; ++i) {
						const float4 *tp = sc.triAccel + (size_t) (first + i) * 3;
						const float4 q0 = __ldg(tp), q1 = __ldg(tp + 1), q2 = __ldg(tp + 2);
						const uint32_t kraw = __float_as_uint(q0.x);
						if (SHADOW && (kraw & TRI_NONOCCLUDER)) continue;
						if (COUNT) triTests++;
						float u, v, t;
						uint32_t nestedUnused = 0xFFFFFFFFu;     // scenes with instances take the lock-step kernels
						if (primTest<SHADOW, SPH>(sc, q0, q1, q2, kraw, s.r, s.mint, s.maxt, u, v, t, nestedUnused)) {
							s.found = true;
							if (SHADOW) break;
							s.maxt = t;                    // sahkdtree3.h:278-283: a later equal t replaces an earlier one
							s.t = t; s.u = u; s.v = v; s.tri = first + i;
						}
					}
					if (SHADOW && s.found) { s.sp = 0; s.cur = BVH_EMPTY; }          // any-hit: done
					else if ((s.cur & BVH_LEAF) && s.cur != BVH_EMPTY) { s.leaf = s.cur; STACK_POP(s.cur); }
				}
			}
			// (C) rays with nothing left to visit are finished
			if (s.active && s.cur == BVH_EMPTY && s.leaf == BVH_EMPTY) {
				s.active = false;
				if (SHADOW) sink.any(s.ray, s.found); else sink.closest(s.ray, s.found, s.t, s.u, s.v, s.tri);
			}
			const uint32_t act = __ballot_sync(0xFFFFFFFFu, s.active);
			if (act == 0) break;
			if (__popc(act) < rbNow && !__all_sync(0xFFFFFFFFu, exhausted || s.active)) break;
		}
	}
#undef STACK_PUSH
#undef STACK_POP
}

// counter builds: per-warp sums of node visits / primitive tests into two 64-bit device counters
__device__ __forceinline__ void flushCounters(uint32_t nodeVisits, uint32_t triTests, unsigned long long *counters) {
	for (int o = 16; o > 0; o >>= 1) {
		nodeVisits += __shfl_xor_sync(0xFFFFFFFFu, nodeVisits, o);
		triTests += __shfl_xor_sync(0xFFFFFFFFu, triTests, o);
	}
	if ((threadIdx.x & 31) == 0) {
		atomicAdd(&counters[0], (unsigned long long) nodeVisits);
		atomicAdd(&counters[1], (unsigned long long) triTests);
	}
}

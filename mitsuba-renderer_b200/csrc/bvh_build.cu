// GPU acceleration-structure build. Replaces ShapeKDTree::build (src/librender/skdtree.cpp:62-106)
// and GenericKDTree::buildInternal (include/mitsuba/render/gkdtree.h:913-1214):
//   1. k_prepare_tris   TriAccel::load for every triangle, bit-exact (triaccel.h:63-96), triangle
//                       boxes and the tight scene box
//   2. k_morton + radix sort of 63-bit Morton keys (the sort itself is cub::DeviceRadixSort)
//   3. k_lbvh_hierarchy binary radix tree over the sorted keys (Karras 2012)
//   4. k_refit          bottom-up boxes with one atomic counter per inner node
//   5. k_collapse       binary tree -> 4-wide nodes of 128 B, widest-surface-area child opened
//                       first; subtrees of <= BVH_LEAF_MAX triangles become leaves
//   6. k_reorder        TriAccel records gathered into leaf order so a leaf is one contiguous run
// Only hit results must agree with the reference's kd-tree, not the tree.
#include "context.h"
#ifndef TRACE_SMEM_STACK
#define TRACE_SMEM_STACK 20
#endif
#define TRACE_STACK_FAST TRACE_SMEM_STACK       // = TRACE_SMEM_STACK, TRACE_SMEM_STACK + TRACE_LOCAL_STACK of trace.cuh (checked there)
#define TRACE_STACK_TOTAL 96
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <algorithm>
#include <cfloat>
#include <cstdlib>
#include <cstring>

namespace {

__device__ __forceinline__ int floatToOrdered(float f) {
	int i = __float_as_int(f);
	return i >= 0 ? i : i ^ 0x7FFFFFFF;
}
inline float orderedToFloat(int i) {
	int j = i >= 0 ? i : i ^ 0x7FFFFFFF;
	float f; memcpy(&f, &j, 4); return f;
}

__device__ __forceinline__ float comp(const float4 &v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }

// TriAccel::load (triaccel.h:63-96) with every operation rounded separately, as the
// reference's SSE build does (no FMA contraction).
__device__ void triAccelLoad(const float4 &A, const float4 &B, const float4 &C, TriAccelRec &r) {
	const float bx = __fsub_rn(C.x, A.x), by = __fsub_rn(C.y, A.y), bz = __fsub_rn(C.z, A.z);
	const float cx = __fsub_rn(B.x, A.x), cy = __fsub_rn(B.y, A.y), cz = __fsub_rn(B.z, A.z);
	// N = cross(c, b)
	float N[3] = {
		__fsub_rn(__fmul_rn(cy, bz), __fmul_rn(cz, by)),
		__fsub_rn(__fmul_rn(cz, bx), __fmul_rn(cx, bz)),
		__fsub_rn(__fmul_rn(cx, by), __fmul_rn(cy, bx)) };
	const float b[3] = { bx, by, bz }, c[3] = { cx, cy, cz };
	int k = 0;
	for (int j = 0; j < 3; ++j)
		if (fabsf(N[j]) > fabsf(N[k])) k = j;
	const int u = (k + 1) % 3, v = (k + 2) % 3;   // waldModulo
	const float n_k = N[k];
	const float denom = __fsub_rn(__fmul_rn(b[u], c[v]), __fmul_rn(b[v], c[u]));
	if (denom == 0) {
		r.k = 3; r.n_u = r.n_v = r.n_d = r.a_u = r.a_v = r.b_nu = r.b_nv = r.c_nu = r.c_nv = 0;
		return;
	}
	r.k = (uint32_t) k;
	r.n_u = __fdiv_rn(N[u], n_k);
	r.n_v = __fdiv_rn(N[v], n_k);
	const float dotAN = __fadd_rn(__fadd_rn(__fmul_rn(A.x, N[0]), __fmul_rn(A.y, N[1])), __fmul_rn(A.z, N[2]));
	r.n_d = __fdiv_rn(dotAN, n_k);
	r.b_nu = __fdiv_rn(b[u], denom);
	r.b_nv = __fdiv_rn(-b[v], denom);
	r.a_u = comp(A, u);
	r.a_v = comp(A, v);
	r.c_nu = __fdiv_rn(c[v], denom);
	r.c_nv = __fdiv_rn(-c[u], denom);
}

__global__ void k_prepare_tris(const uint4 *triIdx, const float4 *pos, const ShapeInfo *shapes, uint32_t n,
		TriAccelRec *accelOrig, float4 *triLo, float4 *triHi, int *bounds) {
	const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
	if (g < n) {
		const uint4 t = triIdx[g];
		const ShapeInfo sh = shapes[t.w];
		TriAccelRec r;
		if (t.y == 0xFFFFFFFFu) {
			// analytic sphere: centre in the vertex, radius in its w; box = centre -+ |radius| (Sphere::getAABB, sphere.cpp:80-86)
			const float4 c = pos[t.x];
			memset(&r, 0, sizeof(r));
			r.k = 3u; r.n_u = c.x; r.n_v = c.y; r.n_d = c.z; r.a_u = c.w;
			r.shapeIndex = t.w; r.primIndex = 0xFFFFFFFFu;           // KNoTriangleFlag (triaccel.h:27, skdtree.cpp:98)
			const float ar = fabsf(c.w);
			lo[0] = __fsub_rn(c.x, ar); lo[1] = __fsub_rn(c.y, ar); lo[2] = __fsub_rn(c.z, ar);
			hi[0] = __fadd_rn(c.x, ar); hi[1] = __fadd_rn(c.y, ar); hi[2] = __fadd_rn(c.z, ar);
		} else if (t.y == 0xFFFFFFFEu) {
			// instance of a shape group: its world box sits in the two vertices (Instance::getAABB); a_v = 1 marks the
			// slot as an instance, b_nu carries the index into the instance table
			const float4 a = pos[t.x], b = pos[t.z];
			memset(&r, 0, sizeof(r));
			r.k = 3u; r.a_v = sh.isCylinder ? 2.0f : 1.0f; r.b_nu = __uint_as_float(sh.cdfOffset);       // a_v = 2: cylinder
			r.shapeIndex = t.w; r.primIndex = 0xFFFFFFFFu;
			lo[0] = a.x; lo[1] = a.y; lo[2] = a.z; hi[0] = b.x; hi[1] = b.y; hi[2] = b.z;
		} else {
			const float4 A = pos[t.x], B = pos[t.y], C = pos[t.z];
			triAccelLoad(A, B, C, r);
			r.shapeIndex = t.w;
			r.primIndex = g - sh.triOffset;
			lo[0] = fminf(A.x, fminf(B.x, C.x)); hi[0] = fmaxf(A.x, fmaxf(B.x, C.x));
			lo[1] = fminf(A.y, fminf(B.y, C.y)); hi[1] = fmaxf(A.y, fmaxf(B.y, C.y));
			lo[2] = fminf(A.z, fminf(B.z, C.z)); hi[2] = fmaxf(A.z, fmaxf(B.z, C.z));
		}
		if (sh.materialId < 0 && !sh.isInstance) r.k |= TRI_NONOCCLUDER;      // a cylinder without a BSDF is no occluder either     // Instance::m_occluder = true (instance.cpp:25)
		accelOrig[g] = r;
		triLo[g] = make_float4(lo[0], lo[1], lo[2], 0);
		triHi[g] = make_float4(hi[0], hi[1], hi[2], 0);
	}
	// warp reduce, then one atomic per warp and component
	for (int c = 0; c < 3; ++c) {
		float l = lo[c], h = hi[c];
		for (int o = 16; o > 0; o >>= 1) {
			l = fminf(l, __shfl_xor_sync(0xFFFFFFFFu, l, o));
			h = fmaxf(h, __shfl_xor_sync(0xFFFFFFFFu, h, o));
		}
		if ((threadIdx.x & 31) == 0) {
			atomicMin(&bounds[c], floatToOrdered(l));
			atomicMax(&bounds[3 + c], floatToOrdered(h));
		}
	}
}

__device__ __forceinline__ uint64_t expandBits21(uint32_t v) {
	uint64_t x = v & 0x1FFFFFu;
	x = (x | x << 32) & 0x1F00000000FFFFull;
	x = (x | x << 16) & 0x1F0000FF0000FFull;
	x = (x | x << 8) & 0x100F00F00F00F00Full;
	x = (x | x << 4) & 0x10C30C30C30C30C3ull;
	x = (x | x << 2) & 0x1249249249249249ull;
	return x;
}

__global__ void k_morton(const float4 *triLo, const float4 *triHi, uint32_t n, float3 bmin, float3 invExt,
		uint64_t *keys, uint32_t *vals) {
	const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= n) return;
	const float4 lo = triLo[g], hi = triHi[g];
	const float cx = ((lo.x + hi.x) * 0.5f - bmin.x) * invExt.x;
	const float cy = ((lo.y + hi.y) * 0.5f - bmin.y) * invExt.y;
	const float cz = ((lo.z + hi.z) * 0.5f - bmin.z) * invExt.z;
	const uint32_t qx = (uint32_t) fminf(fmaxf(cx * 2097152.0f, 0.0f), 2097151.0f);
	const uint32_t qy = (uint32_t) fminf(fmaxf(cy * 2097152.0f, 0.0f), 2097151.0f);
	const uint32_t qz = (uint32_t) fminf(fmaxf(cz * 2097152.0f, 0.0f), 2097151.0f);
	keys[g] = (expandBits21(qx) << 2) | (expandBits21(qy) << 1) | expandBits21(qz);
	vals[g] = g;
}

// Binary radix tree: inner nodes 0..n-2 (root 0), leaves n-1..2n-2 (leaf l at n-1+l).
__device__ __forceinline__ int delta(const uint64_t *keys, int n, int i, int j) {
	if (j < 0 || j >= n) return -1;
	const uint64_t a = keys[i], b = keys[j];
	if (a == b) return 64 + __clz(i ^ j);
	return __clzll((long long) (a ^ b));
}

__global__ void k_lbvh_hierarchy(const uint64_t *keys, int n, int2 *children, int2 *range, int *parent) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n - 1) return;
	const int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
	const int dmin = delta(keys, n, i, i - d);
	int lmax = 2;
	while (delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
	int l = 0;
	for (int t = lmax / 2; t >= 1; t /= 2)
		if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
	const int j = i + l * d;
	const int dnode = delta(keys, n, i, j);
	int s = 0, t = l;
	do {
		t = (t + 1) / 2;
		if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
	} while (t > 1);
	const int gamma = i + s * d + min(d, 0);
	const int first = min(i, j), last = max(i, j);
	const int left = (first == gamma) ? (n - 1 + gamma) : gamma;
	const int right = (last == gamma + 1) ? (n - 1 + gamma + 1) : gamma + 1;
	children[i] = make_int2(left, right);
	range[i] = make_int2(first, last);
	parent[left] = i;
	parent[right] = i;
	if (i == 0) parent[0] = -1;
}

__global__ void k_refit(const uint32_t *sortedIdx, const float4 *triLo, const float4 *triHi, int n, float pad,
		const int2 *children, const int *parent, int *flags, float4 *boxLo, float4 *boxHi) {
	const int l = blockIdx.x * blockDim.x + threadIdx.x;
	if (l >= n) return;
	const uint32_t g = sortedIdx[l];
	float4 lo = triLo[g], hi = triHi[g];
	// Conservative pad: the TriAccel test works on rounded plane/barycentric values, so an accepted
	// hit can lie a few ulps of the coordinate magnitude outside the exact triangle box.
	lo.x -= pad; lo.y -= pad; lo.z -= pad; hi.x += pad; hi.y += pad; hi.z += pad;
	int node = n - 1 + l;
	boxLo[node] = lo; boxHi[node] = hi;
	node = parent[node];
	while (node >= 0) {
		__threadfence();
		if (atomicAdd(&flags[node], 1) == 0) return;   // first arrival leaves; second one has both children
		const int2 c = children[node];
		const float4 al = __ldcg(&boxLo[c.x]), ah = __ldcg(&boxHi[c.x]);
		const float4 bl = __ldcg(&boxLo[c.y]), bh = __ldcg(&boxHi[c.y]);
		boxLo[node] = make_float4(fminf(al.x, bl.x), fminf(al.y, bl.y), fminf(al.z, bl.z), 0);
		boxHi[node] = make_float4(fmaxf(ah.x, bh.x), fmaxf(ah.y, bh.y), fmaxf(ah.z, bh.z), 0);
		node = parent[node];
	}
}

// Subtree sizes of the radix tree (leaves 1, inner nodes from their key range)
__global__ void k_counts_from_range(const int2 *range, int n, int *count) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= 2 * n - 1) return;
	count[i] = i >= n - 1 ? 1 : range[i].y - range[i].x + 1;
}
__device__ __forceinline__ float halfArea(const float4 &lo, const float4 &hi) {
	const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
	return dx * dy + dy * dz + dz * dx;
}

// Queue item: {binary node, index of the 4-wide node to fill, position of the subtree's first triangle in the
// final (depth-first) triangle order}. The order is written here, leaf by leaf: `order[pos]` = original triangle.
__global__ void k_collapse(const uint4 *qIn, uint32_t nIn, uint4 *qOut, uint32_t *qOutCount, uint32_t *nodeCount,
		int n, const int2 *children, const int *count, const uint32_t *sortedIdx, const float4 *boxLo, const float4 *boxHi,
		BVH4Node *nodes, uint32_t *order) {
	const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
	if (w >= nIn) return;
	const uint4 item = qIn[w];
	int ch[4]; uint32_t first[4]; int cnt;
	{
		const int2 c = children[item.x];
		ch[0] = c.x; ch[1] = c.y; cnt = 2;
		first[0] = item.z; first[1] = item.z + (uint32_t) count[c.x];
	}
	while (cnt < 4) {
		int best = -1; float bestArea = -1.0f;
		for (int i = 0; i < cnt; ++i) {
			if (count[ch[i]] <= BVH_LEAF_MAX) continue;
			const float a = halfArea(boxLo[ch[i]], boxHi[ch[i]]);
			if (a > bestArea) { bestArea = a; best = i; }
		}
		if (best < 0) break;
		const int2 c = children[ch[best]];
		ch[best] = c.x; ch[cnt] = c.y;
		first[cnt++] = first[best] + (uint32_t) count[c.x];
	}
	float lo[3][4], hi[3][4]; uint32_t cw[4];
	for (int i = 0; i < 4; ++i) {
		if (i >= cnt) {
			// empty slot: lo = hi = +inf, which no finite ray interval can enter (trace.cuh relies on this)
			lo[0][i] = lo[1][i] = lo[2][i] = INFINITY; hi[0][i] = hi[1][i] = hi[2][i] = INFINITY;
			cw[i] = BVH_EMPTY;
			continue;
		}
		const float4 l = boxLo[ch[i]], h = boxHi[ch[i]];
		lo[0][i] = l.x; lo[1][i] = l.y; lo[2][i] = l.z; hi[0][i] = h.x; hi[1][i] = h.y; hi[2][i] = h.z;
		const int c = count[ch[i]];
		if (c <= BVH_LEAF_MAX) {
			cw[i] = BVH_LEAF | ((uint32_t) (c - 1) << 28) | first[i];
			int stack[8]; int sp = 0; uint32_t k = first[i];          // the subtree's triangles, left to right
			stack[sp++] = ch[i];
			while (sp > 0) {
				const int x = stack[--sp];
				if (x >= n - 1) order[k++] = sortedIdx[x - (n - 1)];
				else { const int2 cc = children[x]; stack[sp++] = cc.y; stack[sp++] = cc.x; }
			}
		} else {
			const uint32_t idx = atomicAdd(nodeCount, 1u);
			cw[i] = idx;
			qOut[atomicAdd(qOutCount, 1u)] = make_uint4((uint32_t) ch[i], idx, first[i], 0u);
		}
	}
	BVH4Node nd;
	nd.lox = make_float4(lo[0][0], lo[0][1], lo[0][2], lo[0][3]);
	nd.loy = make_float4(lo[1][0], lo[1][1], lo[1][2], lo[1][3]);
	nd.loz = make_float4(lo[2][0], lo[2][1], lo[2][2], lo[2][3]);
	nd.hix = make_float4(hi[0][0], hi[0][1], hi[0][2], hi[0][3]);
	nd.hiy = make_float4(hi[1][0], hi[1][1], hi[1][2], hi[1][3]);
	nd.hiz = make_float4(hi[2][0], hi[2][1], hi[2][2], hi[2][3]);
	nd.child = make_uint4(cw[0], cw[1], cw[2], cw[3]);
	nd.pad = make_uint4(0, 0, 0, 0);
	nodes[item.y] = nd;
}

// ---- PLOC: parallel locally-ordered clustering (Meister & Bittner 2018) -------------------------------------------
// Bottom-up agglomeration along the Morton order: every cluster looks `r` neighbours to each side for the partner
// that gives the smallest merged box, mutual choices merge, the survivors are compacted, repeat. Better trees than
// the radix tree (merges follow surface area, not key bits) for a few more milliseconds of build.
// Node numbering: leaves n-1+l as in the radix tree; inner nodes are created upwards from 0, so the root is n-2.
#define PLOC_BLOCK 256
#define PLOC_MAX_RADIUS 32

__global__ void k_ploc_init(const uint32_t *sortedIdx, const float4 *triLo, const float4 *triHi, int n, float pad,
		float4 *boxLo, float4 *boxHi, int *count, int *cluster) {
	const int l = blockIdx.x * blockDim.x + threadIdx.x;
	if (l >= n) return;
	const uint32_t g = sortedIdx[l];
	float4 lo = triLo[g], hi = triHi[g];
	lo.x -= pad; lo.y -= pad; lo.z -= pad; hi.x += pad; hi.y += pad; hi.z += pad;      // same conservative pad as k_refit
	boxLo[n - 1 + l] = lo; boxHi[n - 1 + l] = hi;
	count[n - 1 + l] = 1;
	cluster[l] = n - 1 + l;
}

__global__ void __launch_bounds__(PLOC_BLOCK)
k_ploc_nn(const int *cluster, int m, const float4 *boxLo, const float4 *boxHi, int r, int *nn) {
	__shared__ float slo[3][PLOC_BLOCK + 2 * PLOC_MAX_RADIUS], shi[3][PLOC_BLOCK + 2 * PLOC_MAX_RADIUS];
	const int base = blockIdx.x * PLOC_BLOCK - r;
	for (int t = threadIdx.x; t < PLOC_BLOCK + 2 * r; t += PLOC_BLOCK) {
		const int j = base + t;
		if (j >= 0 && j < m) {
			const int c = cluster[j];
			const float4 lo = boxLo[c], hi = boxHi[c];
			slo[0][t] = lo.x; slo[1][t] = lo.y; slo[2][t] = lo.z; shi[0][t] = hi.x; shi[1][t] = hi.y; shi[2][t] = hi.z;
		}
	}
	__syncthreads();
	const int i = blockIdx.x * PLOC_BLOCK + threadIdx.x;
	if (i >= m) return;
	const int ti = threadIdx.x + r;
	const float lx = slo[0][ti], ly = slo[1][ti], lz = slo[2][ti], hx = shi[0][ti], hy = shi[1][ti], hz = shi[2][ti];
	float bestArea = INFINITY; int best = -1;
	for (int o = -r; o <= r; ++o) {
		const int j = i + o;
		if (o == 0 || j < 0 || j >= m) continue;
		const int tj = ti + o;
		const float dx = fmaxf(hx, shi[0][tj]) - fminf(lx, slo[0][tj]);
		const float dy = fmaxf(hy, shi[1][tj]) - fminf(ly, slo[1][tj]);
		const float dz = fmaxf(hz, shi[2][tj]) - fminf(lz, slo[2][tj]);
		const float a = dx * dy + dy * dz + dz * dx;
		if (a < bestArea) { bestArea = a; best = j; }          // ascending j: ties keep the smaller index
	}
	nn[i] = best;
}

// flags: low word 1 if the cluster survives this round (not the higher-indexed half of a merging pair), high word
// 1 if it is the lower-indexed half of a merging pair (a new node is created in its place)
__global__ void k_ploc_flags(const int *nn, int m, unsigned long long *flags) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= m) return;
	const int j = nn[i];
	const bool mutual = j >= 0 && nn[j] == i;
	const unsigned long long valid = (mutual && i > j) ? 0ull : 1ull, merge = (mutual && i < j) ? 1ull : 0ull;
	flags[i] = valid | (merge << 32);
}

__global__ void k_ploc_apply(const int *cluster, int m, const int *nn, const unsigned long long *flags, const unsigned long long *pos,
		int nextId, int2 *children, float4 *boxLo, float4 *boxHi, int *count, int *clusterOut) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= m) return;
	const unsigned long long f = flags[i];
	if (!(f & 1ull)) return;
	int id = cluster[i];
	if (f >> 32) {
		const int other = cluster[nn[i]];
		const int nid = nextId + (int) (pos[i] >> 32);
		children[nid] = make_int2(id, other);
		const float4 al = boxLo[id], ah = boxHi[id], bl = boxLo[other], bh = boxHi[other];
		boxLo[nid] = make_float4(fminf(al.x, bl.x), fminf(al.y, bl.y), fminf(al.z, bl.z), 0);
		boxHi[nid] = make_float4(fmaxf(ah.x, bh.x), fmaxf(ah.y, bh.y), fmaxf(ah.z, bh.z), 0);
		count[nid] = count[id] + count[other];
		id = nid;
	}
	clusterOut[(uint32_t) (pos[i] & 0xFFFFFFFFull)] = id;
}

// Shading records (types.cuh: DeviceScene::shadeTris) in sorted triangle order + the inverse permutation
__global__ void k_shade_tris(const uint32_t *order, const uint4 *triIdx, const float4 *positions, const float4 *normals,
		const ShapeInfo *shapes, const float *sphereXf, uint32_t n, float4 *out, uint32_t *sortedOf) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const uint32_t g = order[i];                       // kd primitive order (global triangle index)
	sortedOf[g] = i;
	const uint4 idx = triIdx[g];
	float4 *o = out + (size_t) i * 6;
	if (idx.y == 0xFFFFFFFFu) {
		// sphere: {centre, shape} {radius, inverted, -, -} {worldToObject 3x3 rows | objectToWorld 3x3 rows in the .w and the
		// last record}: filled from the side table below
		const float4 c = positions[idx.x];
		const float *xf = sphereXf + (size_t) shapes[idx.w].cdfOffset * 20;       // cdfOffset doubles as the sphere-table index
		o[0] = make_float4(c.x, c.y, c.z, __uint_as_float(idx.w));
		o[1] = make_float4(c.w, xf[18], xf[0], xf[1]);                            // radius, inverted, w2o[0][0..1]
		o[2] = make_float4(xf[2], xf[3], xf[4], xf[5]);                           // w2o[0][2], w2o[1][0..2]
		o[3] = make_float4(xf[6], xf[7], xf[8], xf[9]);                           // w2o[2][0..2], o2w[0][0]
		o[4] = make_float4(xf[10], xf[11], xf[12], xf[13]);                       // o2w[0][1..2], o2w[1][0..1]
		o[5] = make_float4(xf[14], xf[15], xf[16], xf[17]);                       // o2w[1][2], o2w[2][0..2]
		return;
	}
	if (idx.y == 0xFFFFFFFEu) {                        // instance: only the shape index is read
		o[0] = make_float4(0, 0, 0, __uint_as_float(idx.w));
		for (int k = 1; k < 6; ++k) o[k] = make_float4(0, 0, 0, 0);
		return;
	}
	const uint32_t prim = g - shapes[idx.w].triOffset;
	const float4 p0 = positions[idx.x], p1 = positions[idx.y], p2 = positions[idx.z];
	const float4 n0 = normals[idx.x], n1 = normals[idx.y], n2 = normals[idx.z];
	o[0] = make_float4(p0.x, p0.y, p0.z, __uint_as_float(idx.w));
	o[1] = make_float4(p1.x, p1.y, p1.z, __uint_as_float(prim));
	o[2] = make_float4(p2.x, p2.y, p2.z, __uint_as_float(idx.x));
	o[3] = make_float4(n0.x, n0.y, n0.z, __uint_as_float(idx.y));
	o[4] = make_float4(n1.x, n1.y, n1.z, __uint_as_float(idx.z));
	o[5] = make_float4(n2.x, n2.y, n2.z, 0.0f);
}

__global__ void k_reorder(const uint32_t *sortedIdx, const TriAccelRec *orig, uint32_t n, TriAccelRec *sorted) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) sorted[i] = orig[sortedIdx[i]];
}

// 128-byte fp32 node -> the 64-byte node of the traversal kernels (types.cuh: BVH4QNode). Per axis: origin = just below the
// smallest child plane, quantum 2^e with (extent / 2^e) <= 254.5, child planes floor / ceil'ed outwards in fp64 with a
// hundredth of a quantum of slack for the decode's rounding (m * S + C in trace.cuh: 2^-17 of a quantum range).
// Empty slots get an inverted box and the child word of slot 0 (k_collapse fills the slots in order, so slot 0 is real).
__global__ void k_quantize_nodes(const BVH4Node *in, uint32_t n, BVH4QNode *out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const BVH4Node nd = in[i];
	const float lo[3][4] = { { nd.lox.x, nd.lox.y, nd.lox.z, nd.lox.w }, { nd.loy.x, nd.loy.y, nd.loy.z, nd.loy.w }, { nd.loz.x, nd.loz.y, nd.loz.z, nd.loz.w } };
	const float hi[3][4] = { { nd.hix.x, nd.hix.y, nd.hix.z, nd.hix.w }, { nd.hiy.x, nd.hiy.y, nd.hiy.z, nd.hiy.w }, { nd.hiz.x, nd.hiz.y, nd.hiz.z, nd.hiz.w } };
	const uint32_t cw[4] = { nd.child.x, nd.child.y, nd.child.z, nd.child.w };
	float origin[3], scale[3]; uint32_t qlo[3] = { 0, 0, 0 }, qhi[3] = { 0, 0, 0 };
	for (int a = 0; a < 3; ++a) {
		float mn = INFINITY, mx = -INFINITY;
		for (int c = 0; c < 4; ++c)
			if (cw[c] != BVH_EMPTY) { mn = fminf(mn, lo[a][c]); mx = fmaxf(mx, hi[a][c]); }
		if (!(mn <= mx)) { mn = 0.0f; mx = 0.0f; }
		// the grid's origin sits a little below the smallest plane so that plane 0 keeps its outward slack too
		int e = -100;
		const double ext0 = (double) mx - (double) mn;
		if (ext0 > 0.0) { int k; frexp(ext0 / 254.0, &k); e = max(-100, min(100, k)); }
		const float org = __double2float_rd((double) mn - 0.04 * ldexp(1.0, e));
		const double ext = (double) mx - (double) org;
		if (ext > 0.0) { int k; frexp(ext / 254.5, &k); e = max(-100, min(100, k)); }     // 2^k >= ext / 254.5
		const double quantum = ldexp(1.0, e);
		origin[a] = org; scale[a] = (float) ldexp(1.0, e + 15);
		for (int c = 0; c < 4; ++c) {
			uint32_t l = 255u, h = 0u;                     // empty: inverted
			if (cw[c] != BVH_EMPTY) {
				const double fl = floor(((double) lo[a][c] - (double) org) / quantum - 0.01), fh = ceil(((double) hi[a][c] - (double) org) / quantum + 0.01);
				l = (uint32_t) fmin(fmax(fl, 0.0), 255.0); h = (uint32_t) fmin(fmax(fh, 0.0), 255.0);
			}
			qlo[a] |= l << (8 * c); qhi[a] |= h << (8 * c);
		}
	}
	BVH4QNode q;
	q.ox = origin[0]; q.oy = origin[1]; q.oz = origin[2]; q.scaleX = scale[0]; q.scaleY = scale[1]; q.scaleZ = scale[2];
	for (int c = 0; c < 4; ++c) q.child[c] = cw[c] != BVH_EMPTY ? cw[c] : cw[0];     // empty slot: inverted box + a real child's word (trace.cuh: entryOrInf)
	q.qlox = qlo[0]; q.qloy = qlo[1]; q.qloz = qlo[2]; q.qhix = qhi[0]; q.qhiy = qhi[1]; q.qhiz = qhi[2];
	out[i] = q;
}

inline unsigned blocksFor(size_t n, unsigned bs) { return (unsigned) ((n + bs - 1) / bs); }

} // namespace

// sah_build.cu
int buildSahTree(mtsgpu_ctx *ctx, const uint32_t *initialOrder, const float4 *triLo, const float4 *triHi, int n, float pad,
	int2 *children, int *count, float4 *boxLo, float4 *boxHi, uint32_t *order, int *levelsOut);

// Host-side TriMesh::computeTangentSpaceBasis (src/librender/trimesh.cpp:547-662): per-vertex dpdu / dpdv from the
// texture coordinates, accumulated over the triangles in order and averaged; same float operations in the same order.
// `nrm` is updated too: zero normals become (1, 0, 0) there. Only dpdu is kept (the shading frame uses nothing else).
namespace {
struct HV { float x, y, z; };
inline HV hsub(HV a, HV b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
inline HV hadd(HV a, HV b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
inline HV hmul(HV a, float f) { return { a.x * f, a.y * f, a.z * f }; }
inline HV hcross(HV a, HV b) { return { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }
inline float hlen2(HV a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
inline float hlen(HV a) { return std::sqrt(hlen2(a)); }
inline HV hdiv(HV a, float f) { const float r = 1.0f / f; return hmul(a, r); }
inline void hcoordinateSystem(HV a, HV &b, HV &c) {                 // src/libcore/util.cpp:602-611
	if (std::fabs(a.x) > std::fabs(a.y)) {
		const float invLen = 1.0f / std::sqrt(a.x * a.x + a.z * a.z);
		b = { -a.z * invLen, 0.0f, a.x * invLen };
	} else {
		const float invLen = 1.0f / std::sqrt(a.y * a.y + a.z * a.z);
		b = { 0.0f, -a.z * invLen, a.y * invLen };
	}
	c = hcross(a, b);
}
}

static void computeTangents(HostMesh &m, std::vector<float> &dpduOut, std::vector<float> &dpdvOut) {
	const size_t nv = m.nVert;
	std::vector<HV> dpdu(nv, HV{ 0, 0, 0 }), dpdv(nv, HV{ 0, 0, 0 });
	std::vector<uint32_t> sharers(nv, 0u);
	HV *N = reinterpret_cast<HV *>(m.nrm.data());
	const HV *P = reinterpret_cast<const HV *>(m.pos.data());
	for (size_t i = 0; i < nv; ++i)
		if (N[i].x == 0 && N[i].y == 0 && N[i].z == 0) N[i] = { 1.0f, 0.0f, 0.0f };
	for (uint32_t t = 0; t < m.nTri; ++t) {
		const uint32_t i0 = m.tri[3 * t], i1 = m.tri[3 * t + 1], i2 = m.tri[3 * t + 2];
		const HV v0 = P[i0], v1 = P[i1], v2 = P[i2];
		const float u0x = m.uv[2 * i0], u0y = m.uv[2 * i0 + 1];
		const HV dP1 = hsub(v1, v0), dP2 = hsub(v2, v0);
		const float d1x = m.uv[2 * i1] - u0x, d1y = m.uv[2 * i1 + 1] - u0y, d2x = m.uv[2 * i2] - u0x, d2y = m.uv[2 * i2 + 1] - u0y;
		float invDet = 1.0f; const float determinant = d1x * d2y - d1y * d2x;
		if (determinant != 0) invDet = 1.0f / determinant;
		HV du = hmul(hsub(hmul(dP1, d2y), hmul(dP2, d1y)), invDet);
		HV dv = hmul(hadd(hmul(dP1, -d2x), hmul(dP2, d1x)), invDet);
		if (hlen(du) == 0.0f) {
			HV n = hcross(hsub(v1, v0), hsub(v2, v0));
			const float length = hlen(n);
			if (length != 0) {
				n = hdiv(n, length);
				du = hcross(n, dv);
				if (hlen(du) == 0.0f) hcoordinateSystem(n, du, dv);
			}
		}
		if (hlen(dv) == 0.0f) {
			HV n = hcross(hsub(v1, v0), hsub(v2, v0));
			const float length = hlen(n);
			if (length != 0) {
				n = hdiv(n, length);
				dv = hcross(du, n);
				if (hlen(dv) == 0.0f) hcoordinateSystem(n, du, dv);
			}
		}
		dpdu[i0] = hadd(dpdu[i0], du); dpdu[i1] = hadd(dpdu[i1], du); dpdu[i2] = hadd(dpdu[i2], du);
		dpdv[i0] = hadd(dpdv[i0], dv); dpdv[i1] = hadd(dpdv[i1], dv); dpdv[i2] = hadd(dpdv[i2], dv);
		sharers[i0]++; sharers[i1]++; sharers[i2]++;
	}
	dpduOut.resize(3 * nv); dpdvOut.resize(3 * nv);
	for (size_t i = 0; i < nv; ++i) {
		if (hlen2(dpdu[i]) == 0.0f || hlen2(dpdv[i]) == 0.0f) {
			hcoordinateSystem(N[i], dpdu[i], dpdv[i]);
		} else if (sharers[i] > 0) {
			dpdu[i] = hdiv(dpdu[i], (float) sharers[i]);
			dpdv[i] = hdiv(dpdv[i], (float) sharers[i]);
		}
		dpduOut[3 * i] = dpdu[i].x; dpduOut[3 * i + 1] = dpdu[i].y; dpduOut[3 * i + 2] = dpdu[i].z;
		dpdvOut[3 * i] = dpdv[i].x; dpdvOut[3 * i + 1] = dpdv[i].y; dpdvOut[3 * i + 2] = dpdv[i].z;
	}
}

// Host-side DiscretePDF::build (include/mitsuba/core/pdf.h:82-100), float arithmetic in the same order.
static float buildCdf(const std::vector<float> &pdf, std::vector<float> &cdf) {
	cdf.assign(pdf.size() + 1, 0.0f);
	for (size_t i = 1; i < cdf.size(); ++i) {
		volatile float s = cdf[i - 1] + pdf[i - 1];
		cdf[i] = s;
	}
	const float sum = cdf.back();
	for (size_t i = 0; i < pdf.size(); ++i) {
		volatile float q = cdf[i] / sum;
		cdf[i] = q;
	}
	cdf.back() = 1.0f;
	return sum;
}

// Luminaire-selection CDF (Scene::initialize, scene.cpp:320-330): weights are the luminaires'
// samplingWeight when importanceSampleLuminaires is set (the default), else 1 each.
static inline float __uint_as_float_host(uint32_t v) { float f; memcpy(&f, &v, 4); return f; }

int rebuildLuminaireCdf(mtsgpu_ctx *ctx, bool importance) {
	if (ctx->lumImportance == (int) importance) return MTSGPU_OK;
	cudaStream_t st = ctx->stream;
	std::vector<EmitterDev> emitters(ctx->emitters.size());
	std::vector<float> lumPdf(ctx->emitters.size()), lumCdf;
	ctx->scene.background = -1;
	for (size_t e = 0; e < emitters.size(); ++e) {
		memset(&emitters[e], 0, sizeof(EmitterDev));
		const mtsgpu_emitter &src = ctx->emitters[e];
		EmitterDev &em = emitters[e];
		memcpy(em.intensity, src.intensity, 12);
		em.shape = (uint32_t) std::max(ctx->emitterShape[e], 0);
		em.weight = src.samplingWeight;
		em.type = src.type;
		lumPdf[e] = importance ? src.samplingWeight : 1.0f;
		// Scene::getBSphere(): ShapeKDTree::build takes it from the enlarged kd-tree box (skdtree.cpp:68, aabb.cpp:44-47)
		float center[3], radius = 0.0f;
		{
			float d2 = 0.0f;
			for (int c = 0; c < 3; ++c) {
				center[c] = (ctx->scene.aabbMax[c] + ctx->scene.aabbMin[c]) * 0.5f;
				const float d = center[c] - ctx->scene.aabbMax[c];
				d2 = c == 0 ? d * d : d2 + d * d;
			}
			radius = std::sqrt(d2);
		}
		switch (src.type) {
			case MTSGPU_EMITTER_POINT:
				memcpy(em.p, src.p, 12);
				break;
			case MTSGPU_EMITTER_SPOT: {            // SpotLuminaire::configure (spot.cpp:55-61)
				memcpy(em.p, src.p, 24);
				const float beamWidth = src.p[6], cutoffAngle = src.p[7];
				em.p[6] = cutoffAngle;
				em.p[7] = 1.0f / (cutoffAngle - beamWidth);
				em.p[8] = std::cos(beamWidth);
				em.p[9] = std::cos(cutoffAngle);
				break;
			}
			case MTSGPU_EMITTER_COLLIMATED:        // collimated.cpp:29-37
				em.p[0] = src.p[0];
				break;
			case MTSGPU_EMITTER_DIRECTIONAL:       // DirectionalLuminaire::preprocess (directional.cpp:64-71)
				memcpy(em.p, src.p, 16);
				if (src.p[3] == 0.0f) em.p[3] = radius;
				break;
			case MTSGPU_EMITTER_ENVMAP:            // EnvMapLuminaire::preprocess (envmap.cpp:112-126): the same sphere
				em.p[4] = src.p[4];
				em.p[5] = __uint_as_float_host(ctx->envTexture);
				/* fall through */
			case MTSGPU_EMITTER_CONSTANT:          // ConstantLuminaire::preprocess (constant.cpp:48-62)
				memcpy(em.p, src.p, 16);
				if (!(src.p[3] > 0.0f)) {
					memcpy(em.p, center, 12);
					float r = radius * 1.01f;
					if (ctx->haveCamera) {
						const float *M = ctx->camera.cameraToWorld;      // Camera::m_position = cameraToWorld(Point(0,0,0))
						float w = M[15], pos[3] = { M[3], M[7], M[11] };
						if (w != 1.0f) { const float rw = 1.0f / w; for (int c = 0; c < 3; ++c) pos[c] *= rw; }
						const float dx = pos[0] - center[0], dy = pos[1] - center[1], dz = pos[2] - center[2];
						const float dist = std::sqrt(dx * dx + dy * dy + dz * dz);
						if (dist > r) r = dist * 1.01f;              // BSphere::expandBy, then the second enlargement
					}
					em.p[3] = r;
				}
				ctx->scene.background = (int32_t) e;
				break;
			default:
				break;
		}
	}
	float lumSum = 0.0f;
	if (!emitters.empty()) {
		lumSum = buildCdf(lumPdf, lumCdf);
		for (size_t e = 0; e < emitters.size(); ++e) {
			volatile float q = lumPdf[e] / lumSum;
			emitters[e].pdf = q;
		}
	} else {
		lumCdf.assign(1, 0.0f);
	}
	{
		std::vector<float> xf = ctx->emitterXf;
		xf.resize(std::max<size_t>(emitters.size(), 1) * 32, 0.0f);
		CUDA_TRY(ctx, ctx->dEmitterXf.alloc(xf.size() * sizeof(float)));
		CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dEmitterXf.p, xf.data(), xf.size() * sizeof(float), cudaMemcpyHostToDevice, st));
		CUDA_TRY(ctx, cudaStreamSynchronize(st));
		ctx->scene.emitterXf = ctx->dEmitterXf.as<float>();
	}
	CUDA_TRY(ctx, ctx->dEmitters.alloc(std::max<size_t>(emitters.size(), 1) * sizeof(EmitterDev)));
	CUDA_TRY(ctx, ctx->dLumCdf.alloc(lumCdf.size() * sizeof(float)));
	if (!emitters.empty())
		CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dEmitters.p, emitters.data(), emitters.size() * sizeof(EmitterDev), cudaMemcpyHostToDevice, st));
	CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dLumCdf.p, lumCdf.data(), lumCdf.size() * sizeof(float), cudaMemcpyHostToDevice, st));
	CUDA_TRY(ctx, cudaStreamSynchronize(st));
	ctx->scene.emitters = ctx->dEmitters.as<EmitterDev>();
	ctx->scene.lumCdf = ctx->dLumCdf.as<float>();
	ctx->scene.nEmitters = (uint32_t) emitters.size();
	ctx->scene.lumOriginalSum = lumSum;
	ctx->lumImportance = (int) importance;
	return MTSGPU_OK;
}

int buildAccel(mtsgpu_ctx *ctx) {
	REQUIRE(ctx, !ctx->meshes.empty(), MTSGPU_ERR_ARG, "mtsgpu_build_accel: no meshes uploaded");
	cudaStream_t st = ctx->stream;
	const size_t nShapes = ctx->meshes.size();
	size_t nTri = 0, nVert = 0;
	for (auto &m : ctx->meshes) { nTri += m.nTri; nVert += m.nVert; }
	REQUIRE(ctx, nTri > 0, MTSGPU_ERR_ARG, "mtsgpu_build_accel: scene has no triangles");
	REQUIRE(ctx, nTri < (1u << 28), MTSGPU_ERR_LIMIT, "mtsgpu_build_accel: more than 2^28 triangles");

	// ---- flatten meshes (global vertex ids; float4-padded vertices for 128-bit loads) ----
	// (value-initialising std::vectors of this size costs as much as filling them: raw arrays)
	struct RawU4 { uint4 *p; explicit RawU4(size_t n) : p(static_cast<uint4 *>(malloc(std::max<size_t>(n, 1) * sizeof(uint4)))) { } ~RawU4() { free(p); } uint4 *data() { return p; } uint4 &operator[](size_t i) { return p[i]; } };
	struct RawF4 { float4 *p; explicit RawF4(size_t n) : p(static_cast<float4 *>(malloc(std::max<size_t>(n, 1) * sizeof(float4)))) { } ~RawF4() { free(p); } float4 *data() { return p; } float4 &operator[](size_t i) { return p[i]; } };
	RawU4 triIdx(nTri);
	RawF4 pos(nVert), nrm(nVert);
	REQUIRE(ctx, triIdx.p && pos.p && nrm.p, MTSGPU_ERR_ARG, "mtsgpu_build_accel: out of host memory");
	bool anyTexcoords = false;
	for (auto &m : ctx->meshes) anyTexcoords |= !m.uv.empty();
	std::vector<float2> uvs(anyTexcoords ? nVert : 1, make_float2(0, 0));
	// TriMesh::configure (trimesh.cpp:288-290): meshes with an anisotropic BSDF get a tangent space
	std::vector<std::vector<float> > meshTangents(nShapes), meshTangentsV(nShapes);
	bool anyTangents = false;
	for (size_t s = 0; s < nShapes; ++s) {
		HostMesh &m = ctx->meshes[s];
		if (m.materialId < 0 || m.materialId >= (int) ctx->materials.size()) continue;
		const mtsgpu_material &mat = ctx->materials[m.materialId];
		const bool anisotropic = (mat.type & MTSGPU_BSDF_TYPE_MASK) == MTSGPU_BSDF_WARD && mat.p[8] != mat.p[9];
		bool usesRayDifferentials = false;            // a bitmap texture in one of the BSDF's slots (ldrtexture.cpp:243-245)
		for (int k = 0; k < 3; ++k)
			if (mat.texture[k] > 0 && mat.texture[k] <= ctx->textures.size())
				usesRayDifferentials |= ctx->textures[mat.texture[k] - 1].type == MTSGPU_TEXTURE_BITMAP;
		if (!(anisotropic || usesRayDifferentials) || m.sphere || m.cylinder || m.instance) continue;       // analytic shapes bring their own dpdu / dpdv (sphere.cpp:146-161, cylinder.cpp:152-155)
		if (!anisotropic && m.uv.empty()) continue;   // computeTangentSpaceBasis returns false without texture coordinates
		REQUIRE(ctx, !m.uv.empty(), MTSGPU_ERR_ARG, "shape %zu: computeTangentSpace(): texture coordinates are required to generate "
			"tangent vectors. If you want to render with an anisotropic material, please make sure that all associated shapes have "
			"valid texture coordinates.", s);
		if (!m.hasNormals) continue;                  // "Vertex normals are required to compute a tangent space basis!" (warning only)
		computeTangents(m, meshTangents[s], meshTangentsV[s]);
		anyTangents = true;
	}
	std::vector<float4> tangents(anyTangents ? 2 * nVert : 2, make_float4(0, 0, 0, 0));     // {dpdu, dpdv} per vertex
	std::vector<ShapeInfo> shapes(nShapes);
	std::vector<float> areaCdfAll;
	std::vector<int> emitterShape(ctx->emitters.size(), -1);
	size_t tOff = 0, vOff = 0;
	bool anySphere = false;
	std::vector<InstanceDev> instances;
	std::vector<float> sphereXf;      // per sphere: worldToObject 3x3, objectToWorld 3x3, inverted, pad
	for (size_t s = 0; s < nShapes; ++s) {
		const HostMesh &m = ctx->meshes[s];
		ShapeInfo &si = shapes[s];
		memset(&si, 0, sizeof(si));
		si.materialId = m.materialId; si.emitterId = m.emitterId;
		si.triOffset = (uint32_t) tOff; si.vertOffset = (uint32_t) vOff;
		si.nTri = m.nTri; si.nVert = m.nVert; si.hasNormals = m.hasNormals ? 1u : 0u;
		si.hasTexcoords = m.uv.empty() ? 0u : 1u;
		si.hasTangents = meshTangents[s].empty() ? 0u : 1u;
		if (m.cylinder) {
			si.isCylinder = 1u; anySphere = true;
			si.cdfOffset = (uint32_t) instances.size();
			InstanceDev inst; memset(&inst, 0, sizeof(inst));
			memcpy(inst.w2o, m.w2o, 64); memcpy(inst.o2w, m.o2w, 64); inst.radius = m.radius; inst.length = m.length;
			instances.push_back(inst);
		}
		if (m.instance) {
			si.isInstance = 1u; anySphere = true;            // non-triangle primitive: the kernels' SPH variants
			si.cdfOffset = (uint32_t) instances.size();
			InstanceDev inst; memset(&inst, 0, sizeof(inst));
			memcpy(inst.w2o, m.w2o, 64); memcpy(inst.o2w, m.o2w, 64); inst.group = (uint32_t) m.groupId;
			instances.push_back(inst);
		}
		if (m.sphere) {
			si.isSphere = 1u; anySphere = true;
			si.cdfOffset = (uint32_t) (sphereXf.size() / 20);       // index into the sphere transform table
			for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) sphereXf.push_back(m.w2o[4 * r + c]);
			for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) sphereXf.push_back(m.o2w[4 * r + c]);
			sphereXf.push_back(m.inverted ? 1.0f : 0.0f); sphereXf.push_back(0.0f);
		}
		for (uint32_t v = 0; v < (meshTangents[s].empty() ? 0u : m.nVert); ++v)
		{
			tangents[2 * (vOff + v)] = make_float4(meshTangents[s][3*v], meshTangents[s][3*v+1], meshTangents[s][3*v+2], 0);
			tangents[2 * (vOff + v) + 1] = make_float4(meshTangentsV[s][3*v], meshTangentsV[s][3*v+1], meshTangentsV[s][3*v+2], 0);
		}
		for (uint32_t v = 0; v < (m.uv.empty() ? 0u : m.nVert); ++v) uvs[vOff + v] = make_float2(m.uv[2*v], m.uv[2*v+1]);
		REQUIRE(ctx, m.materialId < (int) ctx->materials.size(), MTSGPU_ERR_ARG, "shape %zu: material id %d out of range", s, m.materialId);
		REQUIRE(ctx, m.emitterId < (int) ctx->emitters.size(), MTSGPU_ERR_ARG, "shape %zu: emitter id %d out of range", s, m.emitterId);
		{
			// the two long loops of the scene upload, on all host cores (a 10M-triangle scene spent 0.15 s here on one)
			const float *mp = m.pos.data(), *mn = m.hasNormals ? m.nrm.data() : nullptr;
			const float w = m.sphere ? m.radius : 0.0f;
			const int64_t nv = m.nVert, nt = m.nTri;
			float4 *dp = pos.data() + vOff, *dn = nrm.data() + vOff;
#pragma omp parallel for schedule(static) num_threads(hostThreads()) if (nv > 65536)
			for (int64_t v = 0; v < nv; ++v) {
				dp[v] = make_float4(mp[3*v], mp[3*v+1], mp[3*v+2], w);
				dn[v] = mn ? make_float4(mn[3*v], mn[3*v+1], mn[3*v+2], 0) : make_float4(0, 0, 0, 0);
			}
			uint4 *dt = triIdx.data() + tOff;
			if (m.sphere) { for (int64_t t = 0; t < nt; ++t) dt[t] = make_uint4((uint32_t) vOff, SPHERE_MARK, SPHERE_MARK, (uint32_t) s); }
			else if (m.instance || m.cylinder) { for (int64_t t = 0; t < nt; ++t) dt[t] = make_uint4((uint32_t) vOff, INSTANCE_MARK, (uint32_t) vOff + 1, (uint32_t) s); }
			else {
				const uint32_t *mt = m.tri.data();
				int64_t bad = -1;
#pragma omp parallel for schedule(static) num_threads(hostThreads()) reduction(max : bad) if (nt > 65536)
				for (int64_t t = 0; t < nt; ++t) {
					const uint32_t a = mt[3*t], b = mt[3*t+1], c = mt[3*t+2];
					if (a >= m.nVert || b >= m.nVert || c >= m.nVert) bad = std::max(bad, t);
					dt[t] = make_uint4((uint32_t) vOff + a, (uint32_t) vOff + b, (uint32_t) vOff + c, (uint32_t) s);
				}
				REQUIRE(ctx, bad < 0, MTSGPU_ERR_ARG, "shape %zu: triangle %lld references vertex %u >= %u", s, (long long) bad,
					std::max(mt[3*std::max<int64_t>(bad, 0)], std::max(mt[3*std::max<int64_t>(bad, 0)+1], mt[3*std::max<int64_t>(bad, 0)+2])), m.nVert);
			}
		}
		if (m.emitterId >= 0) {
			REQUIRE(ctx, emitterShape[m.emitterId] < 0, MTSGPU_ERR_ARG, "emitter %d is bound to more than one shape", m.emitterId);
			emitterShape[m.emitterId] = (int) s;
		}
		if (m.emitterId >= 0 && m.sphere) {
			// Sphere::m_invSurfaceArea (sphere.cpp:59); points are drawn by Sphere::sampleSolidAngle, no area CDF
			volatile float a = 4 * 3.14159265358979323846f; a = a * m.radius; a = a * m.radius;
			si.invSurfaceArea = 1 / a;
		} else if (m.emitterId >= 0) {
			// TriMesh::configure: area PDF over Triangle::surfaceArea (trimesh.cpp:266-283, triangle.cpp:49-55)
			std::vector<float> areas(m.nTri), cdf;
			for (uint32_t t = 0; t < m.nTri; ++t) {
				const float *p0 = &m.pos[3 * m.tri[3*t]], *p1 = &m.pos[3 * m.tri[3*t+1]], *p2 = &m.pos[3 * m.tri[3*t+2]];
				volatile float ax = p1[0] - p0[0], ay = p1[1] - p0[1], az = p1[2] - p0[2];
				volatile float bx = p2[0] - p0[0], by = p2[1] - p0[1], bz = p2[2] - p0[2];
				volatile float m0 = ay * bz, m1 = az * by, m2 = az * bx, m3 = ax * bz, m4 = ax * by, m5 = ay * bx;
				volatile float cx = m0 - m1, cy = m2 - m3, cz = m4 - m5;
				volatile float q0 = cx * cx, q1 = cy * cy, q2 = cz * cz;
				volatile float s01 = q0 + q1; volatile float s012 = s01 + q2;
				areas[t] = 0.5f * sqrtf(s012);
			}
			const float total = buildCdf(areas, cdf);
			si.cdfOffset = (uint32_t) areaCdfAll.size();
			si.invSurfaceArea = 1.0f / total;
			areaCdfAll.insert(areaCdfAll.end(), cdf.begin(), cdf.end());
		}
		tOff += m.nTri; vOff += m.nVert;
	}
	// area emitters must be bound to exactly one shape and appear in first-shape order (scene.cpp:656-664)
	int lastShape = -1;
	for (size_t e = 0; e < emitterShape.size(); ++e) {
		if (ctx->emitters[e].type != MTSGPU_EMITTER_AREA) {
			REQUIRE(ctx, emitterShape[e] < 0, MTSGPU_ERR_ARG, "emitter %zu is not an area luminaire but shape %d names it", e, emitterShape[e]);
			continue;
		}
		REQUIRE(ctx, emitterShape[e] >= 0, MTSGPU_ERR_ARG, "emitter %zu is not bound to any shape", e);
		REQUIRE(ctx, emitterShape[e] > lastShape, MTSGPU_ERR_ARG, "emitters must be listed in the order of their shapes");
		lastShape = emitterShape[e];
	}

	ctx->emitterShape = emitterShape;

	std::vector<MaterialDev> mats(std::max<size_t>(ctx->materials.size(), 1));
	memset(mats.data(), 0, mats.size() * sizeof(MaterialDev));
	for (size_t i = 0; i < ctx->materials.size(); ++i) {
		mats[i].type = ctx->materials[i].type;
		memcpy(mats[i].p, ctx->materials[i].p, 64);
		for (int k = 0; k < 3; ++k) {
			const uint32_t t = ctx->materials[i].texture[k];
			REQUIRE(ctx, t <= ctx->textures.size(), MTSGPU_ERR_ARG, "material %zu: texture %u of slot %d was not uploaded", i, t, k);
			const uint32_t bt = mats[i].type & MTSGPU_BSDF_TYPE_MASK;
			const bool slotOk = k == 2 ? (mats[i].type & MTSGPU_BSDF_MASKED) != 0 : k == 0 ? (bt != MTSGPU_BSDF_MIRROR && bt != MTSGPU_BSDF_DIELECTRIC)
				: (bt == MTSGPU_BSDF_MICROFACET || bt == MTSGPU_BSDF_PHONG || bt == MTSGPU_BSDF_ROUGHGLASS || bt == MTSGPU_BSDF_WARD);
			REQUIRE(ctx, t == 0 || slotOk, MTSGPU_ERR_UNSUPPORTED, "material %zu: slot %d of this BSDF is a plain spectrum in the reference and cannot be textured", i, k);
			mats[i].tex[k] = t;
		}
	}
	std::vector<TextureDev> texs(std::max<size_t>(ctx->textures.size(), 1));
	memset(texs.data(), 0, texs.size() * sizeof(TextureDev));
	std::vector<float> texels; std::vector<uint32_t> mipLevels;
	for (size_t i = 0; i < ctx->textures.size(); ++i) {
		texs[i].type = ctx->textures[i].type; memcpy(texs[i].p, ctx->textures[i].p, 48);
		if (texs[i].type != MTSGPU_TEXTURE_BITMAP) continue;
		REQUIRE(ctx, i < ctx->bitmaps.size() && ctx->bitmaps[i].width > 0, MTSGPU_ERR_ARG, "texture %zu: the bitmap's pixels were not uploaded (mtsgpu_upload_bitmap)", i);
		texs[i].pad[0] = (uint32_t) (mipLevels.size() / 4);
		buildMipPyramid(ctx->bitmaps[i], (int) texs[i].p[0], (int) texs[i].p[1], texels, mipLevels);
		texs[i].pad[1] = (uint32_t) (mipLevels.size() / 4) - texs[i].pad[0];
	}
	// ---- environment map: its pyramid joins the texture table; the sampling density of EnvMapLuminaire::configure
	// (envmap.cpp:93-110) is a DiscretePDF over MIP level min(3, levels - 1), luminance x sin(theta)
	std::vector<float> envData(20, 0.0f);
	for (size_t e = 0; e < ctx->emitters.size(); ++e) {
		if (ctx->emitters[e].type != MTSGPU_EMITTER_ENVMAP) continue;
		REQUIRE(ctx, ctx->envEmitter == (int) e && ctx->envBitmap.width > 0, MTSGPU_ERR_ARG, "emitter %zu: the environment map was not uploaded (mtsgpu_upload_envmap)", e);
		TextureDev t; memset(&t, 0, sizeof(t));
		t.type = MTSGPU_TEXTURE_BITMAP; t.p[2] = 8.0f; t.p[8] = t.p[9] = 1.0f;      // EWA, repeat, maxAnisotropy 8 (mipmap.h:57-59)
		t.pad[0] = (uint32_t) (mipLevels.size() / 4);
		buildMipPyramid(ctx->envBitmap, 0, 0, texels, mipLevels);
		t.pad[1] = (uint32_t) (mipLevels.size() / 4) - t.pad[0];
		ctx->envTexture = (uint32_t) texs.size();
		if (ctx->textures.empty()) { texs[0] = t; ctx->envTexture = 0; } else texs.push_back(t);
		const int level = std::min(3, (int) t.pad[1] - 1);
		const uint32_t *L = &mipLevels[4 * (t.pad[0] + level)];
		const int resX = (int) L[1], resY = (int) L[2];
		std::vector<float> pdf((size_t) resX * resY), cdf;
		for (int y = 0; y < resY; ++y) {
			const float sinFactor = std::sin(3.14159265358979323846f * (y + .5f) / resY);
			for (int x = 0; x < resX; ++x) {
				const float *c = &texels[4 * ((size_t) L[0] + x + (size_t) y * resX)];
				volatile float lum = c[0] * 0.212671f; volatile float g = c[1] * 0.715160f; volatile float b = c[2] * 0.072169f;
				lum = lum + g; lum = lum + b;                                   // Spectrum::getLuminance
				pdf[x + (size_t) y * resX] = lum * sinFactor;
			}
		}
		const float sum = buildCdf(pdf, cdf);
		for (int k = 0; k < 9; ++k) { envData[k] = ctx->envW2L[k]; envData[9 + k] = ctx->envL2W[k]; }
		envData[18] = (float) resX; envData[19] = (float) resY;
		envData.insert(envData.end(), cdf.begin(), cdf.end());
		for (float v : pdf) { volatile float q = v / sum; envData.push_back((float) q); }
	}
	if (texels.empty()) texels.assign(4, 0.0f);
	if (mipLevels.empty()) mipLevels.assign(4, 0u);
	float mipLut[128]; mipWeightLut(mipLut);

	// ---- upload ----
	CUDA_TRY(ctx, ctx->dTriIdx.alloc(nTri * sizeof(uint4)));
	CUDA_TRY(ctx, ctx->dPositions.alloc(nVert * sizeof(float4)));
	CUDA_TRY(ctx, ctx->dNormals.alloc(nVert * sizeof(float4)));
	CUDA_TRY(ctx, ctx->dShapes.alloc(nShapes * sizeof(ShapeInfo)));
	CUDA_TRY(ctx, ctx->dMaterials.alloc(mats.size() * sizeof(MaterialDev)));
	CUDA_TRY(ctx, ctx->dAreaCdf.alloc(std::max<size_t>(areaCdfAll.size(), 1) * sizeof(float)));
	CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dTriIdx.p, triIdx.data(), nTri * sizeof(uint4), cudaMemcpyHostToDevice, st));
	CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dPositions.p, pos.data(), nVert * sizeof(float4), cudaMemcpyHostToDevice, st));
	CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dNormals.p, nrm.data(), nVert * sizeof(float4), cudaMemcpyHostToDevice, st));
	CUDA_TRY(ctx, ctx->dTangents.alloc(tangents.size() * sizeof(float4)));
	CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dTangents.p, tangents.data(), tangents.size() * sizeof(float4), cudaMemcpyHostToDevice, st));
	CUDA_TRY(ctx, ctx->dTexcoords.alloc(uvs.size() * sizeof(float2)));
	CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dTexcoords.p, uvs.data(), uvs.size() * sizeof(float2), cudaMemcpyHostToDevice, st));
	CUDA_TRY(ctx, ctx->dTextures.alloc(texs.size() * sizeof(TextureDev)));
	CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dTextures.p, texs.data(), texs.size() * sizeof(TextureDev), cudaMemcpyHostToDevice, st));
	CUDA_TRY(ctx, ctx->dTexels.alloc(texels.size() * sizeof(float)));
	CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dTexels.p, texels.data(), texels.size() * sizeof(float), cudaMemcpyHostToDevice, st));
	CUDA_TRY(ctx, ctx->dMipLevels.alloc(mipLevels.size() * sizeof(uint32_t)));
	CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dMipLevels.p, mipLevels.data(), mipLevels.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
	CUDA_TRY(ctx, ctx->dMipLut.alloc(sizeof(mipLut)));
	CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dMipLut.p, mipLut, sizeof(mipLut), cudaMemcpyHostToDevice, st));
	{
		std::vector<DeviceScene> groupScenes;
		for (mtsgpu_ctx *g : ctx->groups) groupScenes.push_back(g->scene);
		if (instances.empty()) { groupScenes.clear(); }
		CUDA_TRY(ctx, ctx->dInstances.alloc(std::max<size_t>(instances.size(), 1) * sizeof(InstanceDev)));
		CUDA_TRY(ctx, ctx->dGroupScenes.alloc(std::max<size_t>(groupScenes.size(), 1) * sizeof(DeviceScene)));
		if (!instances.empty()) {
			for (const InstanceDev &i : instances)
				REQUIRE(ctx, i.length > 0 || (i.group < groupScenes.size() && ctx->groups[i.group]->built), MTSGPU_ERR_ARG, "instance of shape group %u, which is not built", i.group);
			CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dInstances.p, instances.data(), instances.size() * sizeof(InstanceDev), cudaMemcpyHostToDevice, st));
			CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dGroupScenes.p, groupScenes.data(), groupScenes.size() * sizeof(DeviceScene), cudaMemcpyHostToDevice, st));
			CUDA_TRY(ctx, cudaStreamSynchronize(st));
		}
	}
	ctx->scene.hasInstances = instances.empty() ? 0u : 1u;
	CUDA_TRY(ctx, ctx->dEnvData.alloc(envData.size() * sizeof(float)));
	CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dEnvData.p, envData.data(), envData.size() * sizeof(float), cudaMemcpyHostToDevice, st));
	CUDA_TRY(ctx, ctx->dCamera.alloc(sizeof(CameraDev)));
	CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dShapes.p, shapes.data(), nShapes * sizeof(ShapeInfo), cudaMemcpyHostToDevice, st));
	CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dMaterials.p, mats.data(), mats.size() * sizeof(MaterialDev), cudaMemcpyHostToDevice, st));
	if (!areaCdfAll.empty())
		CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dAreaCdf.p, areaCdfAll.data(), areaCdfAll.size() * sizeof(float), cudaMemcpyHostToDevice, st));
	CUDA_TRY(ctx, cudaStreamSynchronize(st));

	cudaEvent_t e0 = ctx->ev[0], e1 = ctx->ev[1];
	CUDA_TRY(ctx, cudaEventRecord(e0, st));

	const int n = (int) nTri;
	DevBuf dTriLo, dTriHi, dBounds, dKeys, dKeysAlt, dVals, dValsAlt, dTemp, dChildren, dRange, dParent, dFlags,
	       dBoxLo, dBoxHi, dQ0, dQ1, dCounts, dNodesTmp;
	CUDA_TRY(ctx, ctx->dTriAccelOrig.alloc(nTri * sizeof(TriAccelRec)));
	CUDA_TRY(ctx, ctx->dTriAccel.alloc(nTri * sizeof(TriAccelRec)));
	CUDA_TRY(ctx, dTriLo.alloc(nTri * sizeof(float4)));
	CUDA_TRY(ctx, dTriHi.alloc(nTri * sizeof(float4)));
	CUDA_TRY(ctx, dBounds.alloc(6 * sizeof(int)));
	{
		int init[6] = { INT32_MAX, INT32_MAX, INT32_MAX, INT32_MIN, INT32_MIN, INT32_MIN };
		CUDA_TRY(ctx, cudaMemcpyAsync(dBounds.p, init, sizeof(init), cudaMemcpyHostToDevice, st));
	}
	k_prepare_tris<<<blocksFor(nTri, 256), 256, 0, st>>>(ctx->dTriIdx.as<uint4>(), ctx->dPositions.as<float4>(),
		ctx->dShapes.as<ShapeInfo>(), (uint32_t) nTri, ctx->dTriAccelOrig.as<TriAccelRec>(),
		dTriLo.as<float4>(), dTriHi.as<float4>(), dBounds.as<int>());
	CUDA_TRY(ctx, cudaGetLastError());
	int enc[6];
	CUDA_TRY(ctx, cudaMemcpyAsync(enc, dBounds.p, sizeof(enc), cudaMemcpyDeviceToHost, st));
	CUDA_TRY(ctx, cudaStreamSynchronize(st));
	float tmin[3], tmax[3];
	for (int c = 0; c < 3; ++c) { tmin[c] = orderedToFloat(enc[c]); tmax[c] = orderedToFloat(enc[3 + c]); }

	// enlarged box, same two statements as gkdtree.h:1172-1176 (note: max uses the already-moved min)
	DeviceScene &sc = ctx->scene;
	float maxAbs = 0.0f;
	for (int c = 0; c < 3; ++c) {
		volatile float ext = tmax[c] - tmin[c];
		volatile float grow = ext * MTS_EPSILON;
		volatile float grow2 = grow + MTS_EPSILON;
		volatile float nmin = tmin[c] - grow2;
		volatile float ext2 = tmax[c] - nmin;
		volatile float g3 = ext2 * MTS_EPSILON;
		volatile float g4 = g3 + MTS_EPSILON;
		volatile float nmax = tmax[c] + g4;
		sc.aabbMin[c] = nmin; sc.aabbMax[c] = nmax;
		maxAbs = std::max(maxAbs, std::max(std::fabs(tmin[c]), std::fabs(tmax[c])));
	}

	// ---- Morton keys + sort ----
	CUDA_TRY(ctx, dKeys.alloc(nTri * 8)); CUDA_TRY(ctx, dKeysAlt.alloc(nTri * 8));
	CUDA_TRY(ctx, dVals.alloc(nTri * 4)); CUDA_TRY(ctx, dValsAlt.alloc(nTri * 4));
	float3 bmin = make_float3(tmin[0], tmin[1], tmin[2]);
	// Isotropic quantisation: one scale (the largest extent) for all three axes, so that Morton cells are
	// cubes. Per-axis normalisation spends every third key bit on the thin axis of a flat scene (terrain,
	// rooms) and produces slab-shaped nodes with a large surface area.
	const float maxExt = std::max(tmax[0] - tmin[0], std::max(tmax[1] - tmin[1], tmax[2] - tmin[2]));
	const float inv = maxExt > 0.0f ? 1.0f / maxExt : 0.0f;
	float3 invExt = make_float3(inv, inv, inv);
	if (getenv("MTSGPU_MORTON_ANISO"))       // experiment switch: the per-axis normalisation
		invExt = make_float3(tmax[0] > tmin[0] ? 1.0f / (tmax[0] - tmin[0]) : 0.0f,
			tmax[1] > tmin[1] ? 1.0f / (tmax[1] - tmin[1]) : 0.0f, tmax[2] > tmin[2] ? 1.0f / (tmax[2] - tmin[2]) : 0.0f);
	k_morton<<<blocksFor(nTri, 256), 256, 0, st>>>(dTriLo.as<float4>(), dTriHi.as<float4>(), (uint32_t) nTri, bmin, invExt,
		dKeys.as<uint64_t>(), dVals.as<uint32_t>());
	CUDA_TRY(ctx, cudaGetLastError());
	{
		cub::DoubleBuffer<uint64_t> kb(dKeys.as<uint64_t>(), dKeysAlt.as<uint64_t>());
		cub::DoubleBuffer<uint32_t> vb(dVals.as<uint32_t>(), dValsAlt.as<uint32_t>());
		size_t tempBytes = 0;
		CUDA_TRY(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tempBytes, kb, vb, n, 0, 63, st));
		CUDA_TRY(ctx, dTemp.alloc(tempBytes));
		CUDA_TRY(ctx, cub::DeviceRadixSort::SortPairs(dTemp.p, tempBytes, kb, vb, n, 0, 63, st));
		if (kb.Current() != dKeys.as<uint64_t>()) std::swap(dKeys.p, dKeysAlt.p);
		if (vb.Current() != dVals.as<uint32_t>()) std::swap(dVals.p, dValsAlt.p);
	}
	const uint64_t *keys = dKeys.as<uint64_t>();
	const uint32_t *sortedIdx = dVals.as<uint32_t>();      // the SAH build replaces it by its own leaf order

	// ---- binary tree: the radix tree + refit (default), or PLOC (MTSGPU_BUILDER=ploc). Measured on B200: PLOC with
	//      r = 16 buys 2-4 % of traversal time on the 10M-triangle scene and nothing on the 1M-triangle grid (both are
	//      regular meshes, which Morton order already clusters well) for 10x the build time, so it stays an option.
	const char *builderEnv = getenv("MTSGPU_BUILDER");
	const bool usePloc = builderEnv && strcmp(builderEnv, "ploc") == 0 && n > 1;
	const bool useSah = !(builderEnv && (strcmp(builderEnv, "ploc") == 0 || strcmp(builderEnv, "lbvh") == 0)) && n > 1;     // the default
	int plocRadius = getenv("MTSGPU_PLOC_RADIUS") ? atoi(getenv("MTSGPU_PLOC_RADIUS")) : 16;
	plocRadius = std::max(1, std::min(plocRadius, PLOC_MAX_RADIUS));
	int rootNode = 0;
	const size_t nBin = 2 * nTri - 1;
	DevBuf dCount, dOrder;
	CUDA_TRY(ctx, dCount.alloc(nBin * sizeof(int)));
	CUDA_TRY(ctx, dOrder.alloc(nTri * sizeof(uint32_t)));
	CUDA_TRY(ctx, dChildren.alloc(std::max<size_t>(nTri - 1, 1) * sizeof(int2)));
	CUDA_TRY(ctx, dRange.alloc(std::max<size_t>(nTri - 1, 1) * sizeof(int2)));
	CUDA_TRY(ctx, dParent.alloc(nBin * sizeof(int)));
	CUDA_TRY(ctx, dFlags.alloc(std::max<size_t>(nTri - 1, 1) * sizeof(int)));
	CUDA_TRY(ctx, dBoxLo.alloc(nBin * sizeof(float4)));
	CUDA_TRY(ctx, dBoxHi.alloc(nBin * sizeof(float4)));
	const float pad = 4e-6f * std::max(maxAbs, 1e-3f);
	int plocRounds = 0, sahLevels = 0;
	DevBuf dSahOrder;
	if (useSah) {
		// binned-SAH top-down build (sah_build.cu); the Morton order is only its initial primitive order
		CUDA_TRY(ctx, dSahOrder.alloc(nTri * sizeof(uint32_t)));
		const int rc = buildSahTree(ctx, sortedIdx, dTriLo.as<float4>(), dTriHi.as<float4>(), n, pad, dChildren.as<int2>(), dCount.as<int>(),
			dBoxLo.as<float4>(), dBoxHi.as<float4>(), dSahOrder.as<uint32_t>(), &sahLevels);
		if (rc) return rc;
		sortedIdx = dSahOrder.as<uint32_t>();
	} else if (usePloc) {
		DevBuf dClusterA, dClusterB, dNN, dPFlags, dPPos, dScanTemp;
		CUDA_TRY(ctx, dClusterA.alloc(nTri * sizeof(int))); CUDA_TRY(ctx, dClusterB.alloc(nTri * sizeof(int)));
		CUDA_TRY(ctx, dNN.alloc(nTri * sizeof(int)));
		CUDA_TRY(ctx, dPFlags.alloc((nTri + 1) * sizeof(unsigned long long))); CUDA_TRY(ctx, dPPos.alloc((nTri + 1) * sizeof(unsigned long long)));
		size_t scanBytes = 0;
		CUDA_TRY(ctx, cub::DeviceScan::ExclusiveSum(nullptr, scanBytes, dPFlags.as<unsigned long long>(), dPPos.as<unsigned long long>(), n + 1, st));
		CUDA_TRY(ctx, dScanTemp.alloc(scanBytes));
		k_ploc_init<<<blocksFor(nTri, 256), 256, 0, st>>>(sortedIdx, dTriLo.as<float4>(), dTriHi.as<float4>(), n, pad,
			dBoxLo.as<float4>(), dBoxHi.as<float4>(), dCount.as<int>(), dClusterA.as<int>());
		CUDA_TRY(ctx, cudaGetLastError());
		int *cur = dClusterA.as<int>(), *nxt = dClusterB.as<int>();
		int m = n, nextId = 0;
		while (m > 1) {
			k_ploc_nn<<<blocksFor(m, PLOC_BLOCK), PLOC_BLOCK, 0, st>>>(cur, m, dBoxLo.as<float4>(), dBoxHi.as<float4>(), plocRadius, dNN.as<int>());
			k_ploc_flags<<<blocksFor(m, 256), 256, 0, st>>>(dNN.as<int>(), m, dPFlags.as<unsigned long long>());
			CUDA_TRY(ctx, cudaMemsetAsync(dPFlags.as<unsigned long long>() + m, 0, sizeof(unsigned long long), st));   // pos[m] = totals
			size_t tb = dScanTemp.bytes;
			CUDA_TRY(ctx, cub::DeviceScan::ExclusiveSum(dScanTemp.p, tb, dPFlags.as<unsigned long long>(), dPPos.as<unsigned long long>(), m + 1, st));
			k_ploc_apply<<<blocksFor(m, 256), 256, 0, st>>>(cur, m, dNN.as<int>(), dPFlags.as<unsigned long long>(), dPPos.as<unsigned long long>(),
				nextId, dChildren.as<int2>(), dBoxLo.as<float4>(), dBoxHi.as<float4>(), dCount.as<int>(), nxt);
			CUDA_TRY(ctx, cudaGetLastError());
			unsigned long long totals = 0;
			CUDA_TRY(ctx, cudaMemcpyAsync(&totals, dPPos.as<unsigned long long>() + m, sizeof(totals), cudaMemcpyDeviceToHost, st));
			CUDA_TRY(ctx, cudaStreamSynchronize(st));
			const int survivors = (int) (totals & 0xFFFFFFFFull), merges = (int) (totals >> 32);
			REQUIRE(ctx, merges > 0 && survivors == m - merges, MTSGPU_ERR_CUDA, "PLOC made no progress (m=%d, merges=%d)", m, merges);
			nextId += merges; m = survivors;
			std::swap(cur, nxt);
			plocRounds++;
		}
		REQUIRE(ctx, nextId == n - 1, MTSGPU_ERR_CUDA, "PLOC built %d inner nodes for %d triangles", nextId, n);
		rootNode = n - 2;
	} else {
		CUDA_TRY(ctx, cudaMemsetAsync(dFlags.p, 0, dFlags.bytes, st));
		CUDA_TRY(ctx, cudaMemsetAsync(dParent.p, 0xFF, nBin * sizeof(int), st));
		if (n > 1) {
			k_lbvh_hierarchy<<<blocksFor(nTri - 1, 256), 256, 0, st>>>(keys, n, dChildren.as<int2>(), dRange.as<int2>(), dParent.as<int>());
			CUDA_TRY(ctx, cudaGetLastError());
		}
		k_refit<<<blocksFor(nTri, 256), 256, 0, st>>>(sortedIdx, dTriLo.as<float4>(), dTriHi.as<float4>(), n, pad,
			dChildren.as<int2>(), dParent.as<int>(), dFlags.as<int>(), dBoxLo.as<float4>(), dBoxHi.as<float4>());
		k_counts_from_range<<<blocksFor(nBin, 256), 256, 0, st>>>(dRange.as<int2>(), n, dCount.as<int>());
		CUDA_TRY(ctx, cudaGetLastError());
	}
	(void) plocRounds; (void) sahLevels;

	// ---- collapse to 4-wide nodes ----
	const size_t maxNodes = nTri + 1;
	CUDA_TRY(ctx, dNodesTmp.alloc(maxNodes * sizeof(BVH4Node)));
	uint32_t nodeCount = 1; int wideDepth = 1;
	if (n == 1) {
		float4 lo, hi;
		CUDA_TRY(ctx, cudaMemcpyAsync(&lo, dBoxLo.p, sizeof(float4), cudaMemcpyDeviceToHost, st));
		CUDA_TRY(ctx, cudaMemcpyAsync(&hi, dBoxHi.p, sizeof(float4), cudaMemcpyDeviceToHost, st));
		CUDA_TRY(ctx, cudaStreamSynchronize(st));
		BVH4Node nd;
		const float inf = INFINITY;
		nd.lox = make_float4(lo.x, inf, inf, inf); nd.loy = make_float4(lo.y, inf, inf, inf); nd.loz = make_float4(lo.z, inf, inf, inf);
		nd.hix = make_float4(hi.x, inf, inf, inf); nd.hiy = make_float4(hi.y, inf, inf, inf); nd.hiz = make_float4(hi.z, inf, inf, inf);
		nd.child = make_uint4(BVH_LEAF | 0u, BVH_EMPTY, BVH_EMPTY, BVH_EMPTY);
		nd.pad = make_uint4(0, 0, 0, 0);
		CUDA_TRY(ctx, cudaMemcpyAsync(dNodesTmp.p, &nd, sizeof(nd), cudaMemcpyHostToDevice, st));
		CUDA_TRY(ctx, cudaMemcpyAsync(dOrder.p, sortedIdx, sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
	} else {
		const size_t qCap = nTri / 4 + 64;
		CUDA_TRY(ctx, dQ0.alloc(qCap * sizeof(uint4))); CUDA_TRY(ctx, dQ1.alloc(qCap * sizeof(uint4)));
		CUDA_TRY(ctx, dCounts.alloc(2 * sizeof(uint32_t)));
		uint4 rootItem = make_uint4((uint32_t) rootNode, 0u, 0u, 0u);
		CUDA_TRY(ctx, cudaMemcpyAsync(dQ0.p, &rootItem, sizeof(rootItem), cudaMemcpyHostToDevice, st));
		uint32_t counts[2] = { 0u, 1u };   // [queue-out count, node count]
		uint32_t nIn = 1; wideDepth = 0;
		uint4 *qIn = dQ0.as<uint4>(), *qOut = dQ1.as<uint4>();
		while (nIn > 0) {
			wideDepth++;
			counts[0] = 0;
			CUDA_TRY(ctx, cudaMemcpyAsync(dCounts.p, counts, sizeof(counts), cudaMemcpyHostToDevice, st));
			k_collapse<<<blocksFor(nIn, 128), 128, 0, st>>>(qIn, nIn, qOut, dCounts.as<uint32_t>(), dCounts.as<uint32_t>() + 1,
				n, dChildren.as<int2>(), dCount.as<int>(), sortedIdx, dBoxLo.as<float4>(), dBoxHi.as<float4>(), dNodesTmp.as<BVH4Node>(),
				dOrder.as<uint32_t>());
			CUDA_TRY(ctx, cudaGetLastError());
			CUDA_TRY(ctx, cudaMemcpyAsync(counts, dCounts.p, sizeof(counts), cudaMemcpyDeviceToHost, st));
			CUDA_TRY(ctx, cudaStreamSynchronize(st));
			nIn = counts[0];
			REQUIRE(ctx, nIn <= qCap, MTSGPU_ERR_LIMIT, "BVH collapse queue overflow (%u > %zu)", nIn, qCap);
			std::swap(qIn, qOut);
		}
		nodeCount = counts[1];
	}
	CUDA_TRY(ctx, ctx->dNodes.alloc((size_t) nodeCount * sizeof(BVH4QNode)));
	k_quantize_nodes<<<blocksFor(nodeCount, 128), 128, 0, st>>>(dNodesTmp.as<BVH4Node>(), nodeCount, ctx->dNodes.as<BVH4QNode>());
	CUDA_TRY(ctx, cudaGetLastError());
	CUDA_TRY(ctx, ctx->dTraceFlags.alloc(16));
	CUDA_TRY(ctx, cudaMemsetAsync(ctx->dTraceFlags.p, 0, 16, st));
	CUDA_TRY(ctx, ctx->dShadeTris.alloc(nTri * 6 * sizeof(float4)));
	CUDA_TRY(ctx, ctx->dSortedOf.alloc(nTri * sizeof(uint32_t)));
	DevBuf dSphereXf;
	CUDA_TRY(ctx, dSphereXf.alloc(std::max<size_t>(sphereXf.size(), 1) * sizeof(float)));
	if (!sphereXf.empty())
		CUDA_TRY(ctx, cudaMemcpyAsync(dSphereXf.p, sphereXf.data(), sphereXf.size() * sizeof(float), cudaMemcpyHostToDevice, st));
	k_shade_tris<<<blocksFor(nTri, 256), 256, 0, st>>>(dOrder.as<uint32_t>(), ctx->dTriIdx.as<uint4>(), ctx->dPositions.as<float4>(),
		ctx->dNormals.as<float4>(), ctx->dShapes.as<ShapeInfo>(), dSphereXf.as<float>(), (uint32_t) nTri, ctx->dShadeTris.as<float4>(),
		ctx->dSortedOf.as<uint32_t>());
	k_reorder<<<blocksFor(nTri, 256), 256, 0, st>>>(dOrder.as<uint32_t>(), ctx->dTriAccelOrig.as<TriAccelRec>(), (uint32_t) nTri, ctx->dTriAccel.as<TriAccelRec>());
	CUDA_TRY(ctx, cudaGetLastError());
	CUDA_TRY(ctx, cudaEventRecord(e1, st));
	CUDA_TRY(ctx, cudaStreamSynchronize(st));
	float ms = 0; cudaEventElapsedTime(&ms, e0, e1);

	sc.triAccel = ctx->dTriAccel.as<float4>();
	sc.nodes = ctx->dNodes.as<float4>();
	sc.triIdx = ctx->dTriIdx.as<uint4>();
	sc.positions = ctx->dPositions.as<float4>();
	sc.normals = ctx->dNormals.as<float4>();
	sc.texcoords = ctx->dTexcoords.as<float2>();
	sc.tangents = ctx->dTangents.as<float4>();
	sc.shadeTris = ctx->dShadeTris.as<float4>();
	sc.sortedOf = ctx->dSortedOf.as<uint32_t>();
	sc.hasSpheres = anySphere ? 1u : 0u;
	sc.textures = ctx->dTextures.as<TextureDev>();
	sc.texels = ctx->dTexels.as<float4>(); sc.mipLevels = ctx->dMipLevels.as<uint4>(); sc.mipLut = ctx->dMipLut.as<float>();
	sc.camera = ctx->dCamera.as<CameraDev>(); sc.envData = ctx->dEnvData.as<float>();
	sc.instances = ctx->dInstances.as<InstanceDev>(); sc.groups = ctx->dGroupScenes.as<DeviceScene>();
	sc.shapes = ctx->dShapes.as<ShapeInfo>();
	sc.materials = ctx->dMaterials.as<MaterialDev>();
	sc.areaCdf = ctx->dAreaCdf.as<float>();
	sc.nShapes = (uint32_t) nShapes;
	sc.nTris = (uint32_t) nTri;
	sc.rootIsLeaf = 0;
	ctx->accel.triangles = nTri;
	ctx->accel.nodes = nodeCount;
	ctx->accel.nodeBytes = (uint64_t) nodeCount * sizeof(BVH4QNode);
	ctx->accel.depth = (uint32_t) wideDepth;
	ctx->accel.builder = useSah ? 2u : (usePloc ? 1u : 0u);
	ctx->accel.buildLevels = (uint32_t) (useSah ? sahLevels : plocRounds);
	sc.traceFlags = ctx->dTraceFlags.as<uint32_t>(); sc.oneBits = 0x3F800000u;
	{
		// MTSGPU_DEBUG_STACK_ENTRIES shrinks the traversal stacks so that tests can provoke the overflow report
		const int total = getenv("MTSGPU_DEBUG_STACK_ENTRIES") ? std::max(atoi(getenv("MTSGPU_DEBUG_STACK_ENTRIES")), 1) : TRACE_STACK_TOTAL;
		sc.stackTotal = std::min(total, (int) TRACE_STACK_TOTAL); sc.stackFast = std::min(sc.stackTotal, (int) TRACE_STACK_FAST);
	}
	ctx->accel.triBytes = (uint64_t) nTri * sizeof(TriAccelRec);
	ctx->accel.buildMs = ms;
	ctx->built = true;
	ctx->lumImportance = -1;
	return rebuildLuminaireCdf(ctx, true);
}

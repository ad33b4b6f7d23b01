// Top-down binned-SAH build of the binary BVH on the GPU, level by level. Replaces the quality-defining part of
// GenericKDTree::buildInternal (include/mitsuba/render/gkdtree.h:913-1214: the reference evaluates the surface-area
// heuristic over min-max bins, gkdtree.h:1735-1898, then exactly per node): here every level of the tree is one set of
// launches over all primitives.
//
//   per level:  k_sah_prepare   node: centroid bounds -> bin transform, final (padded) box
//               k_sah_bin       primitive: 16 bins x 3 axes of {count, box} per node, integer atomics
//               k_sah_split     node: sweep the bins, take the cheapest of 3 x 15 planes (area_L * n_L + area_R * n_R);
//                               allocate the two children
//               k_sah_flag      primitive: left / right of its node's plane
//               exclusive scan of the flags (cub)
//               k_sah_scatter   primitive: stable partition inside its node's segment; child bounds by atomics;
//                               a child of one primitive is a finished leaf
//
// Output is the binary tree in the numbering k_collapse expects (bvh_build.cu): inner nodes 0 .. n-2 with root 0, leaf l
// (= final position l in the primitive order) at n-1+l; children / count / boxLo / boxHi per node, order[l] = primitive.
// Nodes of <= SAH_SMALL primitives are split at the median position (they end up inside one wide-node leaf anyway).
// Only hit results must agree with the reference's kd-tree, not the tree.
#include "context.h"
#include <cub/device/device_scan.cuh>
#include <climits>
#include <cfloat>

#ifndef SAH_BINS
#define SAH_BINS   16
#endif
#define SAH_SMALL  4
#define SAH_MAX_SAH_LEVEL 64      // below this many levels every split is a median split: bounds the depth

namespace {

__device__ __forceinline__ int fOrd(float f) { const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7FFFFFFF; }
__device__ __forceinline__ float fUnord(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7FFFFFFF); }

struct SahNode {
	int first, cnt;
	int cb[6];                 // centroid bounds (ordered ints): min xyz, max xyz
	int bb[6];                 // primitive bounds
	float cmin[3], scale[3];   // bin(c) = min(SAH_BINS - 1, (int) ((c - cmin) * scale)); scale 0: axis has no extent
	int axis, splitBin, nLeft; // axis -1: split at the median position
	int left, right;
};

struct SahBin { int box[6]; int cnt; };      // one of SAH_BINS x 3 per node of the level

__device__ __forceinline__ void boxAtomics(int *b, const float lo[3], const float hi[3]) {
#pragma unroll
	for (int a = 0; a < 3; ++a) { atomicMin(&b[a], fOrd(lo[a])); atomicMax(&b[3 + a], fOrd(hi[a])); }
}

__device__ __forceinline__ int binOf(float c, float cmin, float scale) {
	const int b = (int) ((c - cmin) * scale);
	return min(max(b, 0), SAH_BINS - 1);
}

__global__ void k_sah_init(const uint32_t *initialOrder, const float4 *triLo, const float4 *triHi, int n, SahNode *nodes,
		uint32_t *prim, int *nodeOf, int *nodeOfAlt, int *count) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const uint32_t g = initialOrder[i];
	prim[i] = g; nodeOf[i] = 0; nodeOfAlt[i] = 0;
	count[n - 1 + i] = 1;
	const float4 lo = triLo[g], hi = triHi[g];
	const float l[3] = { lo.x, lo.y, lo.z }, h[3] = { hi.x, hi.y, hi.z };
	const float c[3] = { (lo.x + hi.x) * 0.5f, (lo.y + hi.y) * 0.5f, (lo.z + hi.z) * 0.5f };
	boxAtomics(nodes[0].bb, l, h);
	boxAtomics(nodes[0].cb, c, c);
}

__global__ void k_sah_prepare(SahNode *nodes, int begin, int end, float pad, float4 *boxLo, float4 *boxHi) {
	const int k = begin + blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= end) return;
	SahNode &nd = nodes[k];
#pragma unroll
	for (int a = 0; a < 3; ++a) {
		const float lo = fUnord(nd.cb[a]), hi = fUnord(nd.cb[3 + a]);
		const float ext = hi - lo;
		nd.cmin[a] = lo;
		nd.scale[a] = (ext > 0.0f && ext < FLT_MAX) ? (float) SAH_BINS * (1.0f - 1e-6f) / ext : 0.0f;
	}
	// same conservative pad as the leaves get (bvh_build.cu: k_refit)
	boxLo[k] = make_float4(fUnord(nd.bb[0]) - pad, fUnord(nd.bb[1]) - pad, fUnord(nd.bb[2]) - pad, 0);
	boxHi[k] = make_float4(fUnord(nd.bb[3]) + pad, fUnord(nd.bb[4]) + pad, fUnord(nd.bb[5]) + pad, 0);
}

__global__ void k_sah_clear_bins(SahBin *bins, size_t nBins) {
	const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= nBins) return;
	SahBin b;
	b.box[0] = b.box[1] = b.box[2] = INT_MAX; b.box[3] = b.box[4] = b.box[5] = INT_MIN; b.cnt = 0;
	bins[i] = b;
}

// Blocks whose positions all lie in ONE node (every block of the upper levels: a node's primitives are contiguous) gather
// their bins in shared memory first and flush one atomic per bin word; the first version sent every primitive's 21 atomics
// straight to the node's 48 global bins, and the eight uppermost levels of the 10M-triangle build took 0.8 s of same-address
// atomics (profiles/r02_launches_c4_64spp.csv). Lower levels, where a block spans many nodes, keep the direct atomics.
#define SAH_BLOCK 256
__global__ void __launch_bounds__(SAH_BLOCK)
k_sah_bin(const SahNode *nodes, int begin, const uint32_t *prim, const int *nodeOf, const float4 *triLo, const float4 *triHi,
		int n, SahBin *bins) {
	__shared__ SahBin sb[3 * SAH_BINS];
	const int i = blockIdx.x * SAH_BLOCK + threadIdx.x;
	const int k = i < n ? nodeOf[i] : -2;
	const int kFirst = nodeOf[min(blockIdx.x * SAH_BLOCK, n - 1)];
	const bool uniform = __syncthreads_and(k == kFirst && k >= 0) != 0;
	if (uniform) {
		const SahNode &nd = nodes[k];
		if (nd.cnt <= SAH_SMALL) return;                   // block-uniform
		for (int t = threadIdx.x; t < 3 * SAH_BINS; t += SAH_BLOCK) {
			sb[t].box[0] = sb[t].box[1] = sb[t].box[2] = INT_MAX; sb[t].box[3] = sb[t].box[4] = sb[t].box[5] = INT_MIN; sb[t].cnt = 0;
		}
		__syncthreads();
		const uint32_t g = prim[i];
		const float4 lo = triLo[g], hi = triHi[g];
		const float l[3] = { lo.x, lo.y, lo.z }, h[3] = { hi.x, hi.y, hi.z };
#pragma unroll
		for (int a = 0; a < 3; ++a) {
			const float sc = nd.scale[a];
			if (sc == 0.0f) continue;
			SahBin *bb = sb + a * SAH_BINS + binOf((l[a] + h[a]) * 0.5f, nd.cmin[a], sc);
			atomicAdd(&bb->cnt, 1);
			boxAtomics(bb->box, l, h);
		}
		__syncthreads();
		SahBin *nb = bins + (size_t) (k - begin) * (3 * SAH_BINS);
		for (int t = threadIdx.x; t < 3 * SAH_BINS; t += SAH_BLOCK) {
			if (sb[t].cnt == 0) continue;
			atomicAdd(&nb[t].cnt, sb[t].cnt);
			for (int d = 0; d < 3; ++d) { atomicMin(&nb[t].box[d], sb[t].box[d]); atomicMax(&nb[t].box[3 + d], sb[t].box[3 + d]); }
		}
		return;
	}
	if (k < 0) return;
	const SahNode &nd = nodes[k];
	if (nd.cnt <= SAH_SMALL) return;
	const uint32_t g = prim[i];
	const float4 lo = triLo[g], hi = triHi[g];
	const float l[3] = { lo.x, lo.y, lo.z }, h[3] = { hi.x, hi.y, hi.z };
	SahBin *nb = bins + (size_t) (k - begin) * (3 * SAH_BINS);
#pragma unroll
	for (int a = 0; a < 3; ++a) {
		const float sc = nd.scale[a];
		if (sc == 0.0f) continue;
		SahBin *b = nb + a * SAH_BINS + binOf((l[a] + h[a]) * 0.5f, nd.cmin[a], sc);
		atomicAdd(&b->cnt, 1);
		boxAtomics(b->box, l, h);
	}
}

__device__ __forceinline__ float areaOf(const float lo[3], const float hi[3]) {
	const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
	return dx * dy + dy * dz + dz * dx;
}

__global__ void k_sah_split(SahNode *nodes, int begin, int end, const SahBin *bins, int n, int forceMedian, int *nodeCounter,
		int2 *children, int *count) {
	const int k = begin + blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= end) return;
	SahNode &nd = nodes[k];
	int bestAxis = -1, bestBin = 0, bestLeft = nd.cnt / 2;
	if (nd.cnt > SAH_SMALL && !forceMedian) {
		float bestCost = INFINITY;
		const SahBin *nb = bins + (size_t) (k - begin) * (3 * SAH_BINS);
		for (int a = 0; a < 3; ++a) {
			if (nd.scale[a] == 0.0f) continue;
			const SahBin *b = nb + a * SAH_BINS;
			float rArea[SAH_BINS]; int rCnt[SAH_BINS];
			{
				float lo[3] = { INFINITY, INFINITY, INFINITY }, hi[3] = { -INFINITY, -INFINITY, -INFINITY }; int c = 0;
				for (int j = SAH_BINS - 1; j >= 1; --j) {
					if (b[j].cnt > 0) {
						for (int d = 0; d < 3; ++d) { lo[d] = fminf(lo[d], fUnord(b[j].box[d])); hi[d] = fmaxf(hi[d], fUnord(b[j].box[3 + d])); }
						c += b[j].cnt;
					}
					rArea[j] = c > 0 ? areaOf(lo, hi) : 0.0f; rCnt[j] = c;
				}
			}
			float lo[3] = { INFINITY, INFINITY, INFINITY }, hi[3] = { -INFINITY, -INFINITY, -INFINITY }; int c = 0;
			for (int j = 1; j < SAH_BINS; ++j) {            // plane j: bins [0, j) left, [j, SAH_BINS) right
				if (b[j - 1].cnt > 0) {
					for (int d = 0; d < 3; ++d) { lo[d] = fminf(lo[d], fUnord(b[j - 1].box[d])); hi[d] = fmaxf(hi[d], fUnord(b[j - 1].box[3 + d])); }
					c += b[j - 1].cnt;
				}
				if (c == 0 || rCnt[j] == 0) continue;
				const float cost = areaOf(lo, hi) * (float) c + rArea[j] * (float) rCnt[j];
				if (cost < bestCost) { bestCost = cost; bestAxis = a; bestBin = j; bestLeft = c; }
			}
		}
		if (bestAxis < 0) bestLeft = nd.cnt / 2;
	}
	nd.axis = bestAxis; nd.splitBin = bestBin; nd.nLeft = bestLeft;
	const int nRight = nd.cnt - bestLeft;
	int ids[2];
	const int cnts[2] = { bestLeft, nRight }, firsts[2] = { nd.first, nd.first + bestLeft };
	for (int s = 0; s < 2; ++s) {
		if (cnts[s] == 1) { ids[s] = n - 1 + firsts[s]; continue; }
		const int id = atomicAdd(nodeCounter, 1);
		ids[s] = id;
		SahNode &ch = nodes[id];
		ch.first = firsts[s]; ch.cnt = cnts[s];
		ch.cb[0] = ch.cb[1] = ch.cb[2] = INT_MAX; ch.cb[3] = ch.cb[4] = ch.cb[5] = INT_MIN;
		ch.bb[0] = ch.bb[1] = ch.bb[2] = INT_MAX; ch.bb[3] = ch.bb[4] = ch.bb[5] = INT_MIN;
		count[id] = cnts[s];
	}
	nd.left = ids[0]; nd.right = ids[1];
	children[k] = make_int2(ids[0], ids[1]);
}

__global__ void k_sah_flag(const SahNode *nodes, const uint32_t *prim, const int *nodeOf, const float4 *triLo, const float4 *triHi, int n, uint32_t *flags) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i > n) return;
	uint32_t f = 0;
	if (i < n) {
		const int k = nodeOf[i];
		if (k >= 0) {
			const SahNode &nd = nodes[k];
			if (nd.axis < 0) f = (i - nd.first) < nd.nLeft ? 1u : 0u;
			else {
				const uint32_t g = prim[i];
				const float4 lo = triLo[g], hi = triHi[g];
				const float c = nd.axis == 0 ? (lo.x + hi.x) * 0.5f : (nd.axis == 1 ? (lo.y + hi.y) * 0.5f : (lo.z + hi.z) * 0.5f);
				f = binOf(c, nd.cmin[nd.axis], nd.scale[nd.axis]) < nd.splitBin ? 1u : 0u;
			}
		}
	}
	flags[i] = f;
}

__global__ void __launch_bounds__(SAH_BLOCK)
k_sah_scatter(SahNode *nodes, const uint32_t *prim, const int *nodeOf, const uint32_t *flags, const uint32_t *scan,
		const float4 *triLo, const float4 *triHi, int n, float pad, uint32_t *primOut, int *nodeOfOut, int *nodeOfIn,
		uint32_t *order, float4 *boxLo, float4 *boxHi) {
	__shared__ int sBounds[2][12];             // [left / right child][bb 6, cb 6] of a block that lies in one node
	const int i = blockIdx.x * SAH_BLOCK + threadIdx.x;
	const int k = i < n ? nodeOf[i] : -2;
	const int kFirst = nodeOf[min(blockIdx.x * SAH_BLOCK, n - 1)];
	const bool uniform = __syncthreads_and(k == kFirst && k >= 0) != 0;
	if (uniform) {
		if (threadIdx.x < 24) sBounds[threadIdx.x / 12][threadIdx.x % 12] = (threadIdx.x % 6) < 3 ? INT_MAX : INT_MIN;
		__syncthreads();
	}
	if (k >= 0) {
		const SahNode &nd = nodes[k];
		const uint32_t g = prim[i];
		const int leftRank = (int) (scan[i] - scan[nd.first]);
		const bool isLeft = flags[i] != 0;
		const int newPos = isLeft ? nd.first + leftRank : nd.first + nd.nLeft + (i - nd.first - leftRank);
		const int child = isLeft ? nd.left : nd.right;
		const int childCnt = isLeft ? nd.nLeft : nd.cnt - nd.nLeft;
		const float4 lo = triLo[g], hi = triHi[g];
		if (childCnt == 1) {
			order[newPos] = g;
			boxLo[n - 1 + newPos] = make_float4(lo.x - pad, lo.y - pad, lo.z - pad, 0);
			boxHi[n - 1 + newPos] = make_float4(hi.x + pad, hi.y + pad, hi.z + pad, 0);
			nodeOfOut[newPos] = -1;          // the other buffer's slot is cleared by k_sah_retire (another thread may still read it)
		} else {
			primOut[newPos] = g; nodeOfOut[newPos] = child;
			const float l[3] = { lo.x, lo.y, lo.z }, h[3] = { hi.x, hi.y, hi.z };
			const float c[3] = { (lo.x + hi.x) * 0.5f, (lo.y + hi.y) * 0.5f, (lo.z + hi.z) * 0.5f };
			int *bb = uniform ? &sBounds[isLeft ? 0 : 1][0] : nodes[child].bb, *cb = uniform ? &sBounds[isLeft ? 0 : 1][6] : nodes[child].cb;
			boxAtomics(bb, l, h);
			boxAtomics(cb, c, c);
		}
	}
	if (uniform) {
		__syncthreads();
		if (threadIdx.x < 24) {
			const SahNode &nd = nodes[kFirst];
			const int side = threadIdx.x / 12, w = threadIdx.x % 12;
			const int child = side == 0 ? nd.left : nd.right, childCnt = side == 0 ? nd.nLeft : nd.cnt - nd.nLeft;
			const int v = sBounds[side][w];
			if (childCnt > 1 && v != ((w % 6) < 3 ? INT_MAX : INT_MIN)) {
				int *dst = w < 6 ? &nodes[child].bb[w] : &nodes[child].cb[w - 6];
				if ((w % 6) < 3) atomicMin(dst, v); else atomicMax(dst, v);
			}
		}
	}
	(void) nodeOfIn;
}

// finished positions must read "finished" in both ping-pong buffers
__global__ void k_sah_retire(const int *nodeOfNew, int *nodeOfOld, int n) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n && nodeOfNew[i] < 0) nodeOfOld[i] = -1;
}

inline unsigned blocksFor(size_t n, unsigned bs) { return (unsigned) ((n + bs - 1) / bs); }

} // namespace

// initialOrder: any permutation of the primitives (the Morton order keeps the gathers of neighbouring positions close).
// Returns the number of levels in *levelsOut. n >= 2.
int buildSahTree(mtsgpu_ctx *ctx, const uint32_t *initialOrder, const float4 *triLo, const float4 *triHi, int n, float pad,
		int2 *children, int *count, float4 *boxLo, float4 *boxHi, uint32_t *order, int *levelsOut) {
	cudaStream_t st = ctx->stream;
	DevBuf dNodes, dPrimA, dPrimB, dNodeOfA, dNodeOfB, dFlags, dScan, dScanTemp, dBins, dCounter;
	CUDA_TRY(ctx, dNodes.alloc((size_t) (n - 1) * sizeof(SahNode)));
	CUDA_TRY(ctx, dPrimA.alloc((size_t) n * 4)); CUDA_TRY(ctx, dPrimB.alloc((size_t) n * 4));
	CUDA_TRY(ctx, dNodeOfA.alloc((size_t) n * 4)); CUDA_TRY(ctx, dNodeOfB.alloc((size_t) n * 4));
	CUDA_TRY(ctx, dFlags.alloc((size_t) (n + 1) * 4)); CUDA_TRY(ctx, dScan.alloc((size_t) (n + 1) * 4));
	CUDA_TRY(ctx, dCounter.alloc(sizeof(int)));
	size_t scanBytes = 0;
	CUDA_TRY(ctx, cub::DeviceScan::ExclusiveSum(nullptr, scanBytes, dFlags.as<uint32_t>(), dScan.as<uint32_t>(), n + 1, st));
	CUDA_TRY(ctx, dScanTemp.alloc(scanBytes));

	SahNode *nodes = dNodes.as<SahNode>();
	{
		SahNode root; memset(&root, 0, sizeof(root));
		root.first = 0; root.cnt = n;
		for (int a = 0; a < 3; ++a) { root.cb[a] = root.bb[a] = INT_MAX; root.cb[3 + a] = root.bb[3 + a] = INT_MIN; }
		CUDA_TRY(ctx, cudaMemcpyAsync(nodes, &root, sizeof(root), cudaMemcpyHostToDevice, st));
		const int one = 1;
		CUDA_TRY(ctx, cudaMemcpyAsync(dCounter.p, &one, sizeof(int), cudaMemcpyHostToDevice, st));
		CUDA_TRY(ctx, cudaMemcpyAsync(count, &n, sizeof(int), cudaMemcpyHostToDevice, st));
	}
	uint32_t *prim = dPrimA.as<uint32_t>(), *primAlt = dPrimB.as<uint32_t>();
	int *nodeOf = dNodeOfA.as<int>(), *nodeOfAlt = dNodeOfB.as<int>();
	k_sah_init<<<blocksFor(n, 256), 256, 0, st>>>(initialOrder, triLo, triHi, n, nodes, prim, nodeOf, nodeOfAlt, count);
	CUDA_TRY(ctx, cudaGetLastError());

	int begin = 0, end = 1, levels = 0;
	while (begin < end) {
		const int width = end - begin;
		REQUIRE(ctx, levels < 4096, MTSGPU_ERR_LIMIT, "SAH build did not terminate (level %d)", levels);
		const size_t nBins = (size_t) width * 3 * SAH_BINS;
		CUDA_TRY(ctx, dBins.alloc(nBins * sizeof(SahBin)));
		k_sah_prepare<<<blocksFor(width, 128), 128, 0, st>>>(nodes, begin, end, pad, boxLo, boxHi);
		k_sah_clear_bins<<<blocksFor(nBins, 256), 256, 0, st>>>(dBins.as<SahBin>(), nBins);
		k_sah_bin<<<blocksFor(n, SAH_BLOCK), SAH_BLOCK, 0, st>>>(nodes, begin, prim, nodeOf, triLo, triHi, n, dBins.as<SahBin>());
		k_sah_split<<<blocksFor(width, 64), 64, 0, st>>>(nodes, begin, end, dBins.as<SahBin>(), n, levels >= SAH_MAX_SAH_LEVEL ? 1 : 0,
			dCounter.as<int>(), children, count);
		k_sah_flag<<<blocksFor((size_t) n + 1, 256), 256, 0, st>>>(nodes, prim, nodeOf, triLo, triHi, n, dFlags.as<uint32_t>());
		CUDA_TRY(ctx, cudaGetLastError());
		size_t tb = dScanTemp.bytes;
		CUDA_TRY(ctx, cub::DeviceScan::ExclusiveSum(dScanTemp.p, tb, dFlags.as<uint32_t>(), dScan.as<uint32_t>(), n + 1, st));
		k_sah_scatter<<<blocksFor(n, SAH_BLOCK), SAH_BLOCK, 0, st>>>(nodes, prim, nodeOf, dFlags.as<uint32_t>(), dScan.as<uint32_t>(), triLo, triHi, n, pad,
			primAlt, nodeOfAlt, nodeOf, order, boxLo, boxHi);
		k_sah_retire<<<blocksFor(n, 256), 256, 0, st>>>(nodeOfAlt, nodeOf, n);
		CUDA_TRY(ctx, cudaGetLastError());
		int total = 0;
		CUDA_TRY(ctx, cudaMemcpyAsync(&total, dCounter.p, sizeof(int), cudaMemcpyDeviceToHost, st));
		CUDA_TRY(ctx, cudaStreamSynchronize(st));
		REQUIRE(ctx, total <= n - 1, MTSGPU_ERR_CUDA, "SAH build created %d inner nodes for %d primitives", total, n);
		begin = end; end = total;
		std::swap(prim, primAlt); std::swap(nodeOf, nodeOfAlt);
		levels++;
	}
	REQUIRE(ctx, end == n - 1, MTSGPU_ERR_CUDA, "SAH build created %d inner nodes for %d primitives", end, n);
	if (levelsOut) *levelsOut = levels;
	return MTSGPU_OK;
}

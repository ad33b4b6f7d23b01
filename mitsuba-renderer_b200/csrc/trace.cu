// Batch ray-query kernels behind mtsgpu_trace / mtsgpu_trace_device and the wavefront loop.
#include "context.h"
#include "trace.cuh"
static_assert(TRACE_SMEM_STACK + TRACE_LOCAL_STACK == 96, "bvh_build.cu: TRACE_STACK_FAST / TRACE_STACK_TOTAL");
#include <algorithm>
#include <cstdlib>

namespace {

// Output sinks of the persistent traversal (trace.cuh)
struct HitSink {
	const DeviceScene &sc; mtsgpu_hit *out;
	__device__ __forceinline__ void closest(uint32_t ray, bool found, float t, float u, float v, uint32_t tri) {
		mtsgpu_hit h; h.t = INFINITY; h.u = h.v = 0; h.shape = h.prim = 0xFFFFFFFFu;
		if (found) {
			const float4 q2 = __ldg(sc.triAccel + (size_t) tri * 3 + 2);
			h.t = t; h.u = u; h.v = v; h.shape = __float_as_uint(q2.z); h.prim = __float_as_uint(q2.w);
		}
		out[ray] = h;
	}
	__device__ __forceinline__ void any(uint32_t, bool) { }
};
struct OccludedSink {
	uint32_t *out;
	__device__ __forceinline__ void closest(uint32_t, bool, float, float, float, uint32_t) { }
	__device__ __forceinline__ void any(uint32_t ray, bool occluded) { out[ray] = occluded ? 1u : 0u; }
};

// rays: 2 x float4 per ray {o.xyz, mint}{d.xyz, maxt}; 128-bit loads. Persistent warps pull rays from
// `rayCounter` (zeroed by the host before the launch).
template <bool COUNT, bool SPH>
__global__ void __launch_bounds__(TRACE_BLOCK, TRACE_PERSISTENT_BLOCKS_PER_SM)
k_trace_closest(const DeviceScene sc, const float4 *__restrict__ rays, uint32_t n, mtsgpu_hit *__restrict__ out,
		uint32_t *rayCounter, unsigned long long *counters, int refetchBelow, int speculate, int leafThreshold) {
	__shared__ uint2 sstack[TRACE_SMEM_STACK * TRACE_BLOCK];
	uint32_t nodeVisits = 0, triTests = 0;
	HitSink sink = { sc, out };
	tracePersistent<false, COUNT, HitSink, SPH>(sc, rays, n, rayCounter, sink, sstack, nodeVisits, triTests, refetchBelow, speculate != 0, leafThreshold);
	if (COUNT) flushCounters(nodeVisits, triTests, counters);
}

template <bool COUNT, bool SPH>
__global__ void __launch_bounds__(TRACE_BLOCK, TRACE_PERSISTENT_BLOCKS_PER_SM)
k_trace_any(const DeviceScene sc, const float4 *__restrict__ rays, uint32_t n, uint32_t *__restrict__ out,
		uint32_t *rayCounter, unsigned long long *counters, int refetchBelow, int speculate, int leafThreshold) {
	__shared__ uint2 sstack[TRACE_SMEM_STACK * TRACE_BLOCK];
	uint32_t nodeVisits = 0, triTests = 0;
	OccludedSink sink = { out };
	tracePersistent<true, COUNT, OccludedSink, SPH>(sc, rays, n, rayCounter, sink, sstack, nodeVisits, triTests, refetchBelow, speculate != 0, leafThreshold);
	if (COUNT) flushCounters(nodeVisits, triTests, counters);
}

} // namespace

// Grid of the persistent kernels: one wave of resident CTAs (SM count x CTAs per SM), fewer for small batches.
unsigned tracePersistentBlocks(mtsgpu_ctx *ctx, size_t n) {
	if (ctx->smCount == 0) {
		int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
		ctx->smCount = sms > 0 ? sms : 148;
	}
	const size_t full = (size_t) ctx->smCount * TRACE_PERSISTENT_BLOCKS_PER_SM;
	const size_t want = (n + TRACE_BLOCK - 1) / TRACE_BLOCK;
	return (unsigned) std::max<size_t>(1, std::min(full, want));
}

// Thresholds of the persistent traversal (trace.cuh); the MTSGPU_* variables are experiment switches. Trees of a few levels
// (the 36-triangle Cornell box: rays finish after two or three steps) run faster with the late pair, refetch below 12 lanes
// and leaf phase at 16 parked leaves: C1 26.5 -> 23.6 ms; everything deeper keeps 28 / 6, tuned on the 1M and 10M scenes.
#define TRACE_TINY_TREE_DEPTH 3
int traceRefetchBelow(const mtsgpu_ctx *ctx) {
	static const int v = getenv("MTSGPU_REFETCH_BELOW") ? atoi(getenv("MTSGPU_REFETCH_BELOW")) : -1;
	return v >= 0 ? v : (ctx->accel.depth <= TRACE_TINY_TREE_DEPTH ? 12 : TRACE_REFETCH_BELOW);
}
int traceLeafThreshold(const mtsgpu_ctx *ctx) {
	static const int v = getenv("MTSGPU_LEAF_THRESHOLD") ? atoi(getenv("MTSGPU_LEAF_THRESHOLD")) : -1;
	return v >= 0 ? v : (ctx->accel.depth <= TRACE_TINY_TREE_DEPTH ? 16 : TRACE_LEAF_THRESHOLD);
}

// A zeroed ray counter for the next persistent launch: a ring of counters so that consecutive launches on
// the stream never share one; each is re-zeroed (async memset) right before use.
int traceRayCounter(mtsgpu_ctx *ctx, uint32_t **out) {
	CUDA_TRY(ctx, ctx->dRayCounters.alloc(64 * sizeof(uint32_t)));
	uint32_t *p = ctx->dRayCounters.as<uint32_t>() + (ctx->rayCounterNext++ & 63);
	CUDA_TRY(ctx, cudaMemsetAsync(p, 0, sizeof(uint32_t), ctx->stream));
	*out = p;
	return MTSGPU_OK;
}

// After the stream has been synchronised: did any traversal kernel of this context (or of its shape groups) run out of
// stack? The reference's kd-tree cannot (its builder bounds the depth at MTS_KD_MAXDEPTH = 48 entries, gkdtree.h:34); a
// BVH has no such bound, so an overflow is reported instead of silently dropping a subtree.
int checkTraceFlags(mtsgpu_ctx *ctx) {
	uint32_t any = 0;
	for (size_t i = 0; i <= ctx->groups.size(); ++i) {
		mtsgpu_ctx *c = i == 0 ? ctx : ctx->groups[i - 1];
		if (!c->dTraceFlags.p) continue;
		uint32_t f = 0;
		CUDA_TRY(ctx, cudaMemcpy(&f, c->dTraceFlags.p, sizeof(f), cudaMemcpyDeviceToHost));
		if (f) CUDA_TRY(ctx, cudaMemset(c->dTraceFlags.p, 0, sizeof(f)));
		any |= f;
	}
	REQUIRE(ctx, !(any & 2u), MTSGPU_ERR_LIMIT, "Lookup depth exceeds the prime number table size! (halton / hammersley sampler: a path drew more than %d values)", 1000);
	REQUIRE(ctx, !(any & 1u), MTSGPU_ERR_LIMIT, "traversal stack overflow: the acceleration structure is %u levels deep and a ray needed more than %d stack entries; "
		"results of this call are incomplete", ctx->accel.depth, TRACE_SMEM_STACK + TRACE_LOCAL_STACK);
	return MTSGPU_OK;
}

int launchTrace(mtsgpu_ctx *ctx, const float4 *dRays, size_t n, int mode, void *dOut, uint64_t *dCounters) {
	REQUIRE(ctx, ctx->built, MTSGPU_ERR_ARG, "trace: call mtsgpu_build_accel first");
	REQUIRE(ctx, n < (1ull << 31), MTSGPU_ERR_LIMIT, "trace: at most 2^31-1 rays per call");
	if (n == 0) return MTSGPU_OK;
	const unsigned blocks = tracePersistentBlocks(ctx, n);
	unsigned long long *cnt = reinterpret_cast<unsigned long long *>(dCounters);
	uint32_t *rayCounter = nullptr;
	{ int rc = traceRayCounter(ctx, &rayCounter); if (rc) return rc; }
	const int rb = traceRefetchBelow(ctx), lt = traceLeafThreshold(ctx);
	static const int spec = getenv("MTSGPU_SPECULATE") ? atoi(getenv("MTSGPU_SPECULATE")) : 1;
	const bool sph = ctx->scene.hasSpheres != 0;
#define LAUNCH(K, C, S, OUT_T) K<C, S><<<blocks, TRACE_BLOCK, 0, ctx->stream>>>(ctx->scene, dRays, (uint32_t) n, (OUT_T *) dOut, rayCounter, (C) ? cnt : nullptr, rb, spec, lt)
	if (mode == MTSGPU_TRACE_CLOSEST) {
		if (cnt) { if (sph) LAUNCH(k_trace_closest, true, true, mtsgpu_hit); else LAUNCH(k_trace_closest, true, false, mtsgpu_hit); }
		else { if (sph) LAUNCH(k_trace_closest, false, true, mtsgpu_hit); else LAUNCH(k_trace_closest, false, false, mtsgpu_hit); }
	} else if (mode == MTSGPU_TRACE_ANY) {
		if (cnt) { if (sph) LAUNCH(k_trace_any, true, true, uint32_t); else LAUNCH(k_trace_any, true, false, uint32_t); }
		else { if (sph) LAUNCH(k_trace_any, false, true, uint32_t); else LAUNCH(k_trace_any, false, false, uint32_t); }
	} else {
		return setError(ctx, MTSGPU_ERR_ARG, "trace: unknown mode %d", mode);
	}
#undef LAUNCH
	CUDA_TRY(ctx, cudaGetLastError());
	return MTSGPU_OK;
}

// Batch ray-query kernels behind mtsgpu_trace / mtsgpu_trace_device and the wavefront loop.
#include "context.h"
#include "trace.cuh"

namespace {

// rays: 2 x float4 per ray {o.xyz, mint}{d.xyz, maxt}. One thread per ray; 128-bit loads.
template <bool COUNT>
__global__ void __launch_bounds__(TRACE_BLOCK)
k_trace_closest(const DeviceScene sc, const float4 *__restrict__ rays, uint32_t n, mtsgpu_hit *__restrict__ out,
		unsigned long long *counters) {
	__shared__ uint2 sstack[TRACE_SMEM_STACK * TRACE_BLOCK];
	const uint32_t i = blockIdx.x * TRACE_BLOCK + threadIdx.x;
	uint32_t nodeVisits = 0, triTests = 0;
	if (i < n) {
		const float4 a = __ldg(rays + 2 * (size_t) i), b = __ldg(rays + 2 * (size_t) i + 1);
		RayIn r = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w };
		mtsgpu_hit h; h.t = INFINITY; h.u = h.v = 0; h.shape = h.prim = 0xFFFFFFFFu;
		float mint, maxt;
		if (clipToScene<false>(sc, r, mint, maxt)) {
			HitOut ho;
			if (traverse<false, COUNT>(sc, r, mint, maxt, sstack, ho, nodeVisits, triTests)) {
				const float4 q2 = __ldg(sc.triAccel + (size_t) ho.tri * 3 + 2);
				h.t = ho.t; h.u = ho.u; h.v = ho.v;
				h.shape = __float_as_uint(q2.z); h.prim = __float_as_uint(q2.w);
			}
		}
		out[i] = h;
	}
	if (COUNT) {
		for (int o = 16; o > 0; o >>= 1) {
			nodeVisits += __shfl_xor_sync(0xFFFFFFFFu, nodeVisits, o);
			triTests += __shfl_xor_sync(0xFFFFFFFFu, triTests, o);
		}
		if ((threadIdx.x & 31) == 0) {
			atomicAdd(&counters[0], (unsigned long long) nodeVisits);
			atomicAdd(&counters[1], (unsigned long long) triTests);
		}
	}
}

template <bool COUNT>
__global__ void __launch_bounds__(TRACE_BLOCK)
k_trace_any(const DeviceScene sc, const float4 *__restrict__ rays, uint32_t n, uint32_t *__restrict__ out,
		unsigned long long *counters) {
	__shared__ uint2 sstack[TRACE_SMEM_STACK * TRACE_BLOCK];
	const uint32_t i = blockIdx.x * TRACE_BLOCK + threadIdx.x;
	uint32_t nodeVisits = 0, triTests = 0;
	if (i < n) {
		const float4 a = __ldg(rays + 2 * (size_t) i), b = __ldg(rays + 2 * (size_t) i + 1);
		RayIn r = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w };
		float mint, maxt; uint32_t occluded = 0;
		if (clipToScene<true>(sc, r, mint, maxt)) {
			HitOut ho;
			occluded = traverse<true, COUNT>(sc, r, mint, maxt, sstack, ho, nodeVisits, triTests) ? 1u : 0u;
		}
		out[i] = occluded;
	}
	if (COUNT) {
		for (int o = 16; o > 0; o >>= 1) {
			nodeVisits += __shfl_xor_sync(0xFFFFFFFFu, nodeVisits, o);
			triTests += __shfl_xor_sync(0xFFFFFFFFu, triTests, o);
		}
		if ((threadIdx.x & 31) == 0) {
			atomicAdd(&counters[0], (unsigned long long) nodeVisits);
			atomicAdd(&counters[1], (unsigned long long) triTests);
		}
	}
}

} // namespace

int launchTrace(mtsgpu_ctx *ctx, const float4 *dRays, size_t n, int mode, void *dOut, uint64_t *dCounters) {
	REQUIRE(ctx, ctx->built, MTSGPU_ERR_ARG, "trace: call mtsgpu_build_accel first");
	REQUIRE(ctx, n < (1ull << 31), MTSGPU_ERR_LIMIT, "trace: at most 2^31-1 rays per call");
	if (n == 0) return MTSGPU_OK;
	const unsigned blocks = (unsigned) ((n + TRACE_BLOCK - 1) / TRACE_BLOCK);
	unsigned long long *cnt = reinterpret_cast<unsigned long long *>(dCounters);
	if (mode == MTSGPU_TRACE_CLOSEST) {
		if (cnt) k_trace_closest<true><<<blocks, TRACE_BLOCK, 0, ctx->stream>>>(ctx->scene, dRays, (uint32_t) n, (mtsgpu_hit *) dOut, cnt);
		else k_trace_closest<false><<<blocks, TRACE_BLOCK, 0, ctx->stream>>>(ctx->scene, dRays, (uint32_t) n, (mtsgpu_hit *) dOut, nullptr);
	} else if (mode == MTSGPU_TRACE_ANY) {
		if (cnt) k_trace_any<true><<<blocks, TRACE_BLOCK, 0, ctx->stream>>>(ctx->scene, dRays, (uint32_t) n, (uint32_t *) dOut, cnt);
		else k_trace_any<false><<<blocks, TRACE_BLOCK, 0, ctx->stream>>>(ctx->scene, dRays, (uint32_t) n, (uint32_t *) dOut, nullptr);
	} else {
		return setError(ctx, MTSGPU_ERR_ARG, "trace: unknown mode %d", mode);
	}
	CUDA_TRY(ctx, cudaGetLastError());
	return MTSGPU_OK;
}

// Several GPUs driven by one process (the `mitsuba` CLI case): every context renders a disjoint sample
// range of every pixel on its own GPU, concurrently; then ONE kernel on the first GPU sums the film
// accumulators by reading the other GPUs' buffers directly through NVLink peer mappings (no staging copy,
// no host round trip) and the result is copied out once.
// Replaces the work-unit distribution over workers and the mutex-guarded merge of
// BlockedRenderProcess::processResult (src/librender/renderproc.cpp:106-160, 123-130).
#include "context.h"
#include <atomic>
#include <chrono>
#include <thread>
#include <vector>
#include <cstring>

#define MULTI_MAX_GPUS 16

namespace {

struct FilmPtrs { const float *p[MULTI_MAX_GPUS]; int n; };

// Progress of the first context's render, handed from its worker thread to the thread that called mtsgpu_render_multi: the
// callback contract (mtsgpu.h) is "called on the calling thread" — Mitsuba's ProgressReporter and RenderQueue want a registered
// Thread, which a worker spawned here is not.
struct ProgressRelay { std::atomic<double> fraction { 0.0 }; };
void relayProgress(void *user, double fraction, const float *) { static_cast<ProgressRelay *>(user)->fraction.store(fraction); }

// out[i] = sum over GPUs of film_g[i]; film_0 is local, the others are peer-mapped (loads cross NVLink).
// 128-bit loads; grid sized to cover the SMs a few times (grid-stride).
__global__ void k_reduce_films_peer(const FilmPtrs films, float *__restrict__ out, size_t n4, size_t n) {
	const size_t stride = (size_t) gridDim.x * blockDim.x;
	for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
		float4 acc = reinterpret_cast<const float4 *>(films.p[0])[i];
		for (int g = 1; g < films.n; ++g) {
			const float4 v = reinterpret_cast<const float4 *>(films.p[g])[i];
			acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
		}
		reinterpret_cast<float4 *>(out)[i] = acc;
	}
	// tail (n not a multiple of 4)
	const size_t t = n4 * 4 + (size_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (t < n) {
		float acc = films.p[0][t];
		for (int g = 1; g < films.n; ++g) acc += films.p[g][t];
		out[t] = acc;
	}
}

} // namespace

extern "C" int mtsgpu_render_multi(mtsgpu_ctx **ctxs, int nCtx, const mtsgpu_render_params *params,
		volatile const int *cancel, float *outFilm, mtsgpu_stats *stats) {
	if (!ctxs || nCtx < 1 || !ctxs[0]) return MTSGPU_ERR_ARG;
	mtsgpu_ctx *c0 = ctxs[0];
	REQUIRE(c0, params && outFilm, MTSGPU_ERR_ARG, "render_multi: null argument");
	REQUIRE(c0, nCtx <= MULTI_MAX_GPUS, MTSGPU_ERR_LIMIT, "render_multi: at most %d contexts", MULTI_MAX_GPUS);
	REQUIRE(c0, c0->haveFilm, MTSGPU_ERR_ARG, "render_multi: call mtsgpu_set_film first");
	for (int i = 0; i < nCtx; ++i) {
		REQUIRE(c0, ctxs[i] && ctxs[i]->haveFilm && ctxs[i]->film.cropW == c0->film.cropW && ctxs[i]->film.cropH == c0->film.cropH,
			MTSGPU_ERR_ARG, "render_multi: context %d has no film or a different film size", i);
		for (int j = 0; j < i; ++j)
			REQUIRE(c0, ctxs[j]->device != ctxs[i]->device, MTSGPU_ERR_ARG, "render_multi: contexts %d and %d share GPU %d", j, i, ctxs[i]->device);
	}
	const size_t nFloats = (size_t) c0->film.cropW * c0->film.cropH * 5;
	const uint32_t lo = params->sppBegin, hi = params->sppEnd, span = hi - lo;

	// ---- render: one host thread per GPU, each on its own sample range ----
	std::vector<int> rcs(nCtx, MTSGPU_OK);
	std::vector<mtsgpu_stats> sts(nCtx);
	std::vector<std::thread> threads;
	// the caller's progress callback (set on the first context) is served from THIS thread; the workers only report a fraction
	const mtsgpu_progress_fn userFn = c0->progressFn; void *const userArg = c0->progressUser;
	const double userInterval = c0->progressInterval; const bool userWantFilm = c0->progressWantFilm;
	ProgressRelay relay;
	std::atomic<int> running { nCtx };
	if (userFn) { c0->progressFn = relayProgress; c0->progressUser = &relay; c0->progressWantFilm = false; }
	for (int i = 0; i < nCtx; ++i) {
		threads.emplace_back([&, i]() {
			struct Done { std::atomic<int> &r; ~Done() { r.fetch_sub(1); } } done { running };
			mtsgpu_ctx *c = ctxs[i];
			mtsgpu_render_params p = *params;
			p.sppBegin = lo + (uint32_t) ((uint64_t) span * i / nCtx);
			p.sppEnd = lo + (uint32_t) ((uint64_t) span * (i + 1) / nCtx);
			cudaError_t e = cudaSetDevice(c->device);
			if (e != cudaSuccess) { rcs[i] = setError(c, MTSGPU_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e)); return; }
			e = c->dFilm.alloc(nFloats * sizeof(float));
			if (e != cudaSuccess) { rcs[i] = setError(c, MTSGPU_ERR_CUDA, "film allocation: %s", cudaGetErrorString(e)); return; }
			rcs[i] = renderDevice(c, &p, cancel, c->dFilm.as<float>(), &sts[i]);
		});
	}
	if (userFn) {
		double shown = 0.0;
		auto last = std::chrono::steady_clock::now();
		while (running.load() > 0) {
			std::this_thread::sleep_for(std::chrono::milliseconds(5));
			const double f = relay.fraction.load();
			const auto now = std::chrono::steady_clock::now();
			if (f > shown && f < 1.0 && std::chrono::duration<double>(now - last).count() >= userInterval) {
				userFn(userArg, f, nullptr);      // partial films stay on their GPUs: previews are a single-GPU feature
				shown = f; last = now;
			}
		}
	}
	for (auto &t : threads) t.join();
	if (userFn) { c0->progressFn = userFn; c0->progressUser = userArg; c0->progressWantFilm = userWantFilm; }
	for (int i = 0; i < nCtx; ++i)
		if (rcs[i] != MTSGPU_OK) {
			if (i != 0) c0->error = ctxs[i]->error;
			return rcs[i];
		}

	// ---- reduce on GPU 0 over peer memory ----
	CUDA_TRY(c0, cudaSetDevice(c0->device));
	FilmPtrs fp; memset(&fp, 0, sizeof(fp));
	fp.n = nCtx; fp.p[0] = c0->dFilm.as<float>();
	DevBuf staging;     // only used for peers that cannot be mapped
	for (int i = 1; i < nCtx; ++i) {
		int can = 0;
		CUDA_TRY(c0, cudaDeviceCanAccessPeer(&can, c0->device, ctxs[i]->device));
		if (can) {
			cudaError_t e = cudaDeviceEnablePeerAccess(ctxs[i]->device, 0);
			if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
			if (e != cudaSuccess) can = 0;
		}
		if (can) {
			fp.p[i] = ctxs[i]->dFilm.as<float>();
		} else {
			// no peer mapping (e.g. different PCIe roots): fall back to a device-to-device copy per peer
			if (!staging.p) CUDA_TRY(c0, staging.alloc((size_t) (nCtx - 1) * nFloats * sizeof(float)));
			float *dst = staging.as<float>() + (size_t) (i - 1) * nFloats;
			CUDA_TRY(c0, cudaMemcpyPeerAsync(dst, c0->device, ctxs[i]->dFilm.p, ctxs[i]->device, nFloats * sizeof(float), c0->stream));
			fp.p[i] = dst;
		}
	}
	DevBuf &dSum = c0->dStageOut;
	CUDA_TRY(c0, dSum.alloc(nFloats * sizeof(float)));
	CUDA_TRY(c0, cudaEventRecord(c0->ev[6], c0->stream));
	if (nCtx > 1) {
		int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c0->device);
		const unsigned blocks = (unsigned) std::min<size_t>((size_t) (sms > 0 ? sms : 148) * 8, (nFloats / 4 + 255) / 256 + 1);
		k_reduce_films_peer<<<blocks, 256, 0, c0->stream>>>(fp, dSum.as<float>(), nFloats / 4, nFloats);
		CUDA_TRY(c0, cudaGetLastError());
	}
	CUDA_TRY(c0, cudaEventRecord(c0->ev[7], c0->stream));
	CUDA_TRY(c0, cudaMemcpyAsync(outFilm, nCtx > 1 ? dSum.p : c0->dFilm.p, nFloats * sizeof(float), cudaMemcpyDeviceToHost, c0->stream));
	CUDA_TRY(c0, cudaStreamSynchronize(c0->stream));
	if (stats) {
		mtsgpu_stats S; memset(&S, 0, sizeof(S));
		for (int i = 0; i < nCtx; ++i) {
			S.cameraRays += sts[i].cameraRays; S.normalRays += sts[i].normalRays; S.shadowRays += sts[i].shadowRays;
			S.samples += sts[i].samples; S.kernelLaunches += sts[i].kernelLaunches;
			S.traceClosestMs = std::max(S.traceClosestMs, sts[i].traceClosestMs); S.traceAnyMs = std::max(S.traceAnyMs, sts[i].traceAnyMs);
			S.shadeMs = std::max(S.shadeMs, sts[i].shadeMs); S.totalMs = std::max(S.totalMs, sts[i].totalMs);
		}
		float ms = 0; cudaEventElapsedTime(&ms, c0->ev[6], c0->ev[7]);
		S.totalMs += ms; S.kernelLaunches += nCtx > 1 ? 1 : 0;
		*stats = S;
	}
	if (userFn) userFn(userArg, 1.0, nullptr);
	return MTSGPU_OK;
}

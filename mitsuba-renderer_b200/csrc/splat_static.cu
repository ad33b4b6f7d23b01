// mtsgpu_film_splat_static — the film splat for callers that must not allocate.
//
// Who this is for: code that splats samples while it is not allowed to touch the CUDA allocator — a
// progress/preview refresh running on a latency-critical thread while a render already owns most of
// the GPU's memory (cudaMalloc can fail there, and it synchronises the whole device, stalling the
// render), or a host application that hands the library a fixed memory budget up front and audits
// that nothing grows afterwards. The regular mtsgpu_film_splat stages the samples, the filter table
// and the film accumulators in buffers it gets from cudaMalloc; this variant keeps all three in one
// fixed block of device memory that is part of the library image.
//
// Cost: every GPU that loads libmtscuda.so holds MTSGPU_STATIC_SCRATCH_BYTES (4 MiB) until the process
// exits, whether or not this entry point is ever called.
#include "context.h"
#include <cstring>
#include <cmath>

#define STATIC_MAX_DEVICES 64

// The block itself. In host code this name is only a handle for the runtime's symbol table, not a
// device address; the address is looked up once per device below and handed to the kernel.
static __device__ __align__(256) unsigned char g_splatScratch[MTSGPU_STATIC_SCRATCH_BYTES];

static void *g_scratchAddr[STATIC_MAX_DEVICES];   // per device: resolved address, or null before set-up

namespace {

// Same arithmetic as k_splat in integrate.cu (ImageBlock::putSample, imageblock.h:80-138), one thread per
// sample, accumulating straight into the film block in global memory instead of a shared-memory tile.
__global__ void k_splat_static(const FilmDev film, const float *__restrict__ samples, uint32_t n,
		const float *__restrict__ table, float *filmAcc) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const float *s = samples + (size_t) i * 6;
	const float r = s[2], g = s[3], b = s[4], a = s[5];
	if (r != r || g != g || b != b || r < 0.0f || g < 0.0f || b < 0.0f) return;   // !spec.isValid(): dropped
	// a full-frame block: offset = crop offset, border as for render blocks
	const int border = film.border;
	const int sideX = film.cropW + 2 * border, sideY = film.cropH + 2 * border;
	const float fx = s[0] - 0.5f - (float) (film.cropX - border);
	const float fy = s[1] - 0.5f - (float) (film.cropY - border);
	int xStart = (int) ceilf(fx - film.filterSizeX), xEnd = (int) floorf(fx + film.filterSizeX);
	int yStart = (int) ceilf(fy - film.filterSizeY), yEnd = (int) floorf(fy + film.filterSizeY);
	xStart = max(0, xStart); yStart = max(0, yStart);
	xEnd = min(xEnd, sideX - 1); yEnd = min(yEnd, sideY - 1);
	for (int y = yStart; y <= yEnd; ++y) {
		const int gy = y - border;
		if (gy < 0 || gy >= film.cropH) continue;          // Film::putImageBlock skips rows outside the crop
		const int iy = min((int) (film.factorY * fabsf((float) y - fy)), 15);
		for (int x = xStart; x <= xEnd; ++x) {
			const int gx = x - border;
			if (gx < 0 || gx >= film.cropW) continue;
			const int ix = min((int) (film.factorX * fabsf((float) x - fx)), 15);
			const float w = table[iy * 16 + ix];
			float *dst = filmAcc + ((size_t) gy * film.cropW + gx) * 5;
			atomicAdd(dst + 0, r * w); atomicAdd(dst + 1, g * w); atomicAdd(dst + 2, b * w);
			atomicAdd(dst + 3, a * w); atomicAdd(dst + 4, w);
		}
	}
}

} // namespace

extern "C" int mtsgpu_film_splat_static(mtsgpu_ctx *ctx, const float *samples, size_t n, float *outFilm) {
	if (!ctx) return MTSGPU_ERR_ARG;
	{
		cudaError_t e = cudaSetDevice(ctx->device);
		if (e != cudaSuccess) return setError(ctx, MTSGPU_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
	}
	REQUIRE(ctx, ctx->haveFilm, MTSGPU_ERR_ARG, "film_splat_static: call mtsgpu_set_film first");
	REQUIRE(ctx, outFilm && (n == 0 || samples), MTSGPU_ERR_ARG, "film_splat_static: null buffer");
	REQUIRE(ctx, ctx->device >= 0 && ctx->device < STATIC_MAX_DEVICES, MTSGPU_ERR_LIMIT,
		"film_splat_static: device ordinal %d is beyond the %d devices this entry point tracks", ctx->device, STATIC_MAX_DEVICES);
	const FilmDev film = ctx->film;

	// layout of the block: [filter table 1 KiB][film accumulators][samples]; every piece 256-byte aligned
	const size_t tableBytes = 256 * sizeof(float);
	const size_t filmBytes = (size_t) film.cropW * film.cropH * 5 * sizeof(float);
	const size_t sampleBytes = n * 6 * sizeof(float);
	const size_t filmOff = tableBytes;
	const size_t sampleOff = (filmOff + filmBytes + 255) & ~(size_t) 255;
	const size_t need = sampleOff + sampleBytes;
	REQUIRE(ctx, need <= (size_t) MTSGPU_STATIC_SCRATCH_BYTES, MTSGPU_ERR_LIMIT,
		"film_splat_static: a %dx%d film with %zu samples needs %zu bytes of scratch but the built-in block holds %zu "
		"(1024 + 20 bytes per pixel + 24 bytes per sample); use mtsgpu_film_splat, or split the samples over several calls",
		film.cropW, film.cropH, n, need, (size_t) MTSGPU_STATIC_SCRATCH_BYTES);

	// One-time set-up per device, with that device current: resolve the block's device address.
	if (!g_scratchAddr[ctx->device]) {
		void *p = nullptr;
		cudaError_t e = cudaGetSymbolAddress(&p, g_splatScratch);
		if (e != cudaSuccess || !p)
			return setError(ctx, MTSGPU_ERR_CUDA, "film_splat_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
		g_scratchAddr[ctx->device] = p;
	}
	unsigned char *base = static_cast<unsigned char *>(g_scratchAddr[ctx->device]);
	float *dTable = reinterpret_cast<float *>(base);
	float *dFilm = reinterpret_cast<float *>(base + filmOff);
	float *dSamples = reinterpret_cast<float *>(base + sampleOff);
	cudaStream_t st = ctx->stream;

	CUDA_TRY(ctx, cudaMemcpyAsync(dTable, ctx->filterTable, tableBytes, cudaMemcpyHostToDevice, st));
	// the block still holds whatever the previous call left there: the accumulators must start from zero
	CUDA_TRY(ctx, cudaMemsetAsync(dFilm, 0, filmBytes, st));
	if (n) {
		CUDA_TRY(ctx, cudaMemcpyAsync(dSamples, samples, sampleBytes, cudaMemcpyHostToDevice, st));
		k_splat_static<<<(unsigned) ((n + 127) / 128), 128, 0, st>>>(film, dSamples, (uint32_t) n, dTable, dFilm);
		CUDA_TRY(ctx, cudaGetLastError());
	}
	CUDA_TRY(ctx, cudaMemcpyAsync(outFilm, dFilm, filmBytes, cudaMemcpyDeviceToHost, st));
	CUDA_TRY(ctx, cudaStreamSynchronize(st));
	return MTSGPU_OK;
}

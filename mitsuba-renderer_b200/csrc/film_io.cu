// Film development and image writers: what follows the hot path.
//   EXRFilm::develop        src/films/exrfilm.cpp:156-210   accumulators -> RGBA (normalise by the weight, clamp)
//   Bitmap::saveEXR         src/libcore/bitmap.cpp:599-628  half RGBA scanline file; Imf::Header(w, h) defaults
//   MFilm::develop          src/films/mfilm.cpp:174-224     MATLAB text: luminance (or the spectrum) per pixel
// The normalisation runs on the GPU on the device film (k_develop / k_develop_m); the file writers are host code
// (zlib for the EXR ZIP blocks). OpenEXR itself is not used: the file layout is written out directly.
#include "context.h"
#include <zlib.h>
#include <cstring>
#include <cstdio>
#include <vector>
#include <string>
#include <algorithm>

void setGlobalError(const char *fmt, ...);

namespace {

inline unsigned blocksFor(size_t n, unsigned bs) { return (unsigned) ((n + bs - 1) / bs); }

// exrfilm.cpp:163-176 per pixel; accumulators are {spec.rgb, alpha, weight}
__global__ void k_develop(const float *__restrict__ acc, size_t nPix, int hasAlpha, float4 *__restrict__ rgba) {
	const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= nPix) return;
	const float *p = acc + i * 5;
	float invWeight = 1.0f;
	if (p[4] != 0.0f) invWeight = 1.0f / p[4];
	rgba[i] = make_float4(fmaxf(0.0f, p[0] * invWeight), fmaxf(0.0f, p[1] * invWeight), fmaxf(0.0f, p[2] * invWeight),
		hasAlpha ? p[3] * invWeight : 1.0f);
}

// mfilm.cpp:186-210 per pixel: invWeight * spectrum, or invWeight * luminance (spectrum.h:387-389)
__global__ void k_develop_m(const float *__restrict__ acc, size_t nPix, int spectra, float *__restrict__ out) {
	const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= nPix) return;
	const float *p = acc + i * 5;
	const float invWeight = p[4] > 0 ? 1 / p[4] : 1;
	if (spectra) {
		out[3 * i] = __fmul_rn(invWeight, p[0]); out[3 * i + 1] = __fmul_rn(invWeight, p[1]); out[3 * i + 2] = __fmul_rn(invWeight, p[2]);
	} else {
		const float lum = __fadd_rn(__fadd_rn(__fmul_rn(p[0], 0.212671f), __fmul_rn(p[1], 0.715160f)), __fmul_rn(p[2], 0.072169f));
		out[i] = __fmul_rn(invWeight, lum);
	}
}

// float -> half, round to nearest even (what Imf's `half` does on assignment)
inline uint16_t floatToHalf(float f) {
	uint32_t x; memcpy(&x, &f, 4);
	const uint32_t sign = (x >> 16) & 0x8000u;
	x &= 0x7FFFFFFFu;
	if (x >= 0x7F800000u) return (uint16_t) (sign | 0x7C00u | ((x > 0x7F800000u) ? (0x0200u | ((x >> 13) & 0x03FFu)) : 0u));   // inf / nan
	if (x >= 0x477FF000u) return (uint16_t) (sign | 0x7C00u);                     // rounds to >= 65520: overflow to inf
	if (x < 0x33000001u) return (uint16_t) sign;                                   // below half of the smallest subnormal
	if (x < 0x38800000u) {                                                         // subnormal half
		const int shift = 126 - (int) (x >> 23);                                  // 14..24
		const uint32_t mant = (x & 0x007FFFFFu) | 0x00800000u;
		uint32_t h = mant >> shift;
		const uint32_t rem = mant & ((1u << shift) - 1u), halfway = 1u << (shift - 1);
		if (rem > halfway || (rem == halfway && (h & 1u))) h++;
		return (uint16_t) (sign | h);
	}
	uint32_t h = ((x - 0x38000000u) >> 13);
	const uint32_t rem = x & 0x1FFFu;
	if (rem > 0x1000u || (rem == 0x1000u && (h & 1u))) h++;
	return (uint16_t) (sign | h);
}

struct Out {
	std::vector<unsigned char> b;
	void raw(const void *p, size_t n) { const unsigned char *c = static_cast<const unsigned char *>(p); b.insert(b.end(), c, c + n); }
	void i32(int32_t v) { raw(&v, 4); }
	void f32(float v) { raw(&v, 4); }
	void str(const char *s) { raw(s, strlen(s) + 1); }
	void attr(const char *name, const char *type, int32_t size) { str(name); str(type); i32(size); }
};

int writeFile(const char *path, const void *data, size_t n, const char *mode) {
	FILE *f = fopen(path, mode);
	if (!f) { setGlobalError("cannot create \"%s\"", path); return MTSGPU_ERR_ARG; }
	const size_t put = fwrite(data, 1, n, f);
	fclose(f);
	if (put != n) { setGlobalError("short write on \"%s\"", path); return MTSGPU_ERR_ARG; }
	return MTSGPU_OK;
}

} // namespace

extern "C" {

int mtsgpu_write_exr(const char *path, const float *rgba, int width, int height) {
	if (!path || !rgba || width <= 0 || height <= 0) return MTSGPU_ERR_ARG;
	Out o;
	o.i32(20000630); o.i32(2);                                  // magic, version 2, no flags (scanlines, short names)
	// channels A, B, G, R (alphabetical), HALF, no subsampling
	{
		Out c;
		const char *names[4] = { "A", "B", "G", "R" };
		for (int k = 0; k < 4; ++k) { c.str(names[k]); c.i32(1); const unsigned char z[4] = { 0, 0, 0, 0 }; c.raw(z, 4); c.i32(1); c.i32(1); }
		const unsigned char end = 0; c.raw(&end, 1);
		o.attr("channels", "chlist", (int32_t) c.b.size()); o.raw(c.b.data(), c.b.size());
	}
	{ o.attr("compression", "compression", 1); const unsigned char zip = 3; o.raw(&zip, 1); }              // ZIP_COMPRESSION, 16 lines per block
	{ o.attr("dataWindow", "box2i", 16); o.i32(0); o.i32(0); o.i32(width - 1); o.i32(height - 1); }
	{ o.attr("displayWindow", "box2i", 16); o.i32(0); o.i32(0); o.i32(width - 1); o.i32(height - 1); }
	{ o.attr("lineOrder", "lineOrder", 1); const unsigned char inc = 0; o.raw(&inc, 1); }                  // INCREASING_Y
	{ o.attr("pixelAspectRatio", "float", 4); o.f32(1.0f); }
	{ o.attr("screenWindowCenter", "v2f", 8); o.f32(0.0f); o.f32(0.0f); }
	{ o.attr("screenWindowWidth", "float", 4); o.f32(1.0f); }
	{ const unsigned char end = 0; o.raw(&end, 1); }

	const int linesPerBlock = 16, nBlocks = (height + linesPerBlock - 1) / linesPerBlock;
	const size_t tableAt = o.b.size();
	o.b.resize(tableAt + (size_t) nBlocks * 8);
	const int order[4] = { 3, 2, 1, 0 };                         // A, B, G, R taken from {r, g, b, a}
	std::vector<unsigned char> raw, tmp, packed;
	for (int blk = 0; blk < nBlocks; ++blk) {
		const int y0 = blk * linesPerBlock, y1 = std::min(height, y0 + linesPerBlock);
		raw.clear();
		for (int y = y0; y < y1; ++y)
			for (int k = 0; k < 4; ++k)
				for (int x = 0; x < width; ++x) {
					const uint16_t h = floatToHalf(rgba[((size_t) y * width + x) * 4 + order[k]]);
					raw.push_back((unsigned char) (h & 0xFF)); raw.push_back((unsigned char) (h >> 8));
				}
		// ZIP block: split even / odd bytes, delta-encode, deflate; stored raw when that is not smaller
		const size_t n = raw.size();
		tmp.resize(n);
		{
			size_t a = 0, b = (n + 1) / 2;
			for (size_t i = 0; i < n; ++i) { if (i & 1) tmp[b++] = raw[i]; else tmp[a++] = raw[i]; }
			int prev = tmp[0];
			for (size_t i = 1; i < n; ++i) { const int cur = tmp[i]; tmp[i] = (unsigned char) (cur - prev + 128 + 256); prev = cur; }
		}
		uLongf bound = compressBound((uLong) n);
		packed.resize(bound);
		if (compress(packed.data(), &bound, tmp.data(), (uLong) n) != Z_OK) { setGlobalError("write_exr: deflate failed"); return MTSGPU_ERR_ARG; }
		const uint64_t off = o.b.size();
		memcpy(o.b.data() + tableAt + (size_t) blk * 8, &off, 8);
		o.i32(y0);
		if (bound < n) { o.i32((int32_t) bound); o.raw(packed.data(), bound); }
		else { o.i32((int32_t) n); o.raw(raw.data(), n); }
	}
	return writeFile(path, o.b.data(), o.b.size(), "wb");
}

int mtsgpu_write_mfilm(const char *path, const float *values, int width, int height, int channels) {
	if (!path || !values || width <= 0 || height <= 0 || (channels != 1 && channels != 3)) return MTSGPU_ERR_ARG;
	std::string s = "[";
	char buf[64];
	size_t pos = 0;
	for (int y = 0; y < height; ++y) {
		for (int x = 0; x < width; ++x) {
			for (int c = 0; c < channels; ++c) {
				snprintf(buf, sizeof(buf), "%f", values[pos * channels + c]);
				s += buf;
				if (c + 1 < channels) s += " ";
			}
			if (x + 1 < width) s += ", ";
			++pos;
		}
		if (y + 1 < height) s += ";\n ";
	}
	s += "]\n";
	return writeFile(path, s.data(), s.size(), "w");
}

} // extern "C"

// ---- entry points that take the film accumulators (host or device) ------------------------------------------
static int developCommon(mtsgpu_ctx *ctx, const float *accum, bool deviceIn, int mode, int flag, void *out, bool deviceOut, size_t outFloatsPerPixel) {
	REQUIRE(ctx, ctx->haveFilm, MTSGPU_ERR_ARG, "film_develop: call mtsgpu_set_film first");
	REQUIRE(ctx, accum && out, MTSGPU_ERR_ARG, "film_develop: null buffer");
	const size_t nPix = (size_t) ctx->film.cropW * ctx->film.cropH;
	const float *dAcc = accum;
	if (!deviceIn) {
		CUDA_TRY(ctx, ctx->dStage.alloc(nPix * 5 * sizeof(float)));
		CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dStage.p, accum, nPix * 5 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
		dAcc = ctx->dStage.as<float>();
	}
	void *dOut = out;
	if (!deviceOut) {
		CUDA_TRY(ctx, ctx->dStageOut.alloc(nPix * outFloatsPerPixel * sizeof(float)));
		dOut = ctx->dStageOut.p;
	}
	if (mode == 0) k_develop<<<blocksFor(nPix, 256), 256, 0, ctx->stream>>>(dAcc, nPix, flag, static_cast<float4 *>(dOut));
	else k_develop_m<<<blocksFor(nPix, 256), 256, 0, ctx->stream>>>(dAcc, nPix, flag, static_cast<float *>(dOut));
	CUDA_TRY(ctx, cudaGetLastError());
	if (!deviceOut)
		CUDA_TRY(ctx, cudaMemcpyAsync(out, dOut, nPix * outFloatsPerPixel * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
	CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	return MTSGPU_OK;
}

extern "C" {

int mtsgpu_film_develop(mtsgpu_ctx *ctx, const float *film, int hasAlpha, float *rgba) {
	if (!ctx) return MTSGPU_ERR_ARG;
	cudaSetDevice(ctx->device);
	return developCommon(ctx, film, false, 0, hasAlpha, rgba, false, 4);
}

int mtsgpu_film_develop_device(mtsgpu_ctx *ctx, const float *dFilm, int hasAlpha, float *dRgba) {
	if (!ctx) return MTSGPU_ERR_ARG;
	cudaSetDevice(ctx->device);
	return developCommon(ctx, dFilm, true, 0, hasAlpha, dRgba, true, 4);
}

int mtsgpu_film_save(mtsgpu_ctx *ctx, const float *film, int filmOnDevice, int format, int option, const char *path) {
	if (!ctx || !path) return MTSGPU_ERR_ARG;
	cudaSetDevice(ctx->device);
	REQUIRE(ctx, ctx->haveFilm, MTSGPU_ERR_ARG, "film_save: call mtsgpu_set_film first");
	const int w = ctx->film.cropW, h = ctx->film.cropH;
	const size_t nPix = (size_t) w * h;
	if (format == MTSGPU_FILM_EXR) {
		std::vector<float> rgba(nPix * 4);
		int rc = developCommon(ctx, film, filmOnDevice != 0, 0, option, rgba.data(), false, 4);
		if (rc) return rc;
		rc = mtsgpu_write_exr(path, rgba.data(), w, h);
		return rc ? setError(ctx, rc, "%s", mtsgpu_last_error(nullptr)) : MTSGPU_OK;
	}
	if (format == MTSGPU_FILM_M) {
		const int channels = option ? 3 : 1;
		std::vector<float> vals(nPix * channels);
		int rc = developCommon(ctx, film, filmOnDevice != 0, 1, option, vals.data(), false, channels);
		if (rc) return rc;
		rc = mtsgpu_write_mfilm(path, vals.data(), w, h, channels);
		return rc ? setError(ctx, rc, "%s", mtsgpu_last_error(nullptr)) : MTSGPU_OK;
	}
	return setError(ctx, MTSGPU_ERR_ARG, "film_save: unknown format %d", format);
}

} // extern "C"

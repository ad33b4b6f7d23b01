// Mip-map pyramid of a bitmap texture, host code. Restates MIPMap::MIPMap
// (src/librender/mipmap.cpp:31-137) in fp32, operation by operation, so that every texel the GPU filters is the
// texel the reference's MIPMap holds:
//   * a texture whose width or height is not a power of two is first up-sampled with a 4-tap Lanczos filter
//     (resampleWeights, mipmap.cpp:193-218; lanczosSinc, src/libcore/util.cpp:664-674), X pass then Y pass —
//     including the X pass's use of the image HEIGHT in its bounds test (mipmap.cpp:52), which decides which taps of
//     a non-square image wrap;
//   * the level count comes from the ORIGINAL size (mipmap.cpp:99-102), so an up-sampled texture ends one level
//     short of 1 x 1;
//   * each level is the 2 x 2 box average of the one below, read through getTexel (mipmap.cpp:220-243), whose
//     bounds test treats column 0 as outside (x <= 0): with the `black` / `white` wrap modes column 0 reads as
//     the border colour.
#include "context.h"
#include <cmath>
#include <algorithm>

namespace {

struct C3 { float r, g, b; };
inline C3 add(C3 a, C3 b) { return { a.r + b.r, a.g + b.g, a.b + b.b }; }
inline C3 mul(C3 a, float s) { return { a.r * s, a.g * s, a.b * s }; }

const float kPi = 3.14159265358979323846f, kEps = 1e-4f;

inline int modulo(int a, int b) { const int r = a - int(a / b) * b; return r < 0 ? r + b : r; }      // util.cpp:424-427
inline int clampi(int v, int lo, int hi) { return std::min(std::max(v, lo), hi); }
inline bool isPow2(int i) { return i > 0 && (i & (i - 1)) == 0; }
inline uint32_t roundToPow2(uint32_t i) { i--; i |= i >> 1; i |= i >> 2; i |= i >> 4; i |= i >> 8; i |= i >> 16; return i + 1; }
inline int log2i(uint32_t v) { int r = 0; while ((v >> r) != 0) r++; return r - 1; }

float lanczosSinc(float t, float tau = 2.0f) {
	t = std::abs(t);
	if (t < kEps) return 1.0f;
	else if (t > 1.0f) return 0.0f;
	t *= kPi;
	const float sincTerm = std::sin(t * tau) / (t * tau);
	const float windowTerm = std::sin(t) / t;
	return sincTerm * windowTerm;
}

struct ResampleWeight { int firstTexel; float weight[4]; };

std::vector<ResampleWeight> resampleWeights(int oldRes, int newRes) {
	const float filterWidth = 2.0f;
	std::vector<ResampleWeight> w(newRes);
	for (int i = 0; i < newRes; ++i) {
		const float center = (i + .5f) * oldRes / newRes;
		w[i].firstTexel = (int) std::floor(center - filterWidth + 0.5f);
		float weightSum = 0;
		for (int j = 0; j < 4; ++j) {
			const float pos = w[i].firstTexel + j + .5f;
			const float weight = lanczosSinc((pos - center) / filterWidth);
			weightSum += weight;
			w[i].weight[j] = weight;
		}
		const float invWeights = 1.0f / weightSum;
		for (int j = 0; j < 4; ++j) w[i].weight[j] *= invWeights;
	}
	return w;
}

enum { WRAP_REPEAT = 0, WRAP_BLACK = 1, WRAP_WHITE = 2, WRAP_CLAMP = 3, FILTER_NONE = 2 };

C3 getTexel(const std::vector<C3> &lvl, int levelWidth, int levelHeight, int wrapMode, int x, int y) {
	if (x <= 0 || y < 0 || x >= levelWidth || y >= levelHeight) {
		switch (wrapMode) {
			case WRAP_REPEAT: x = modulo(x, levelWidth); y = modulo(y, levelHeight); break;
			case WRAP_CLAMP: x = clampi(x, 0, levelWidth - 1); y = clampi(y, 0, levelHeight - 1); break;
			case WRAP_BLACK: return { 0.0f, 0.0f, 0.0f };
			case WRAP_WHITE: return { 1.0f, 1.0f, 1.0f };
		}
	}
	return lvl[x + (size_t) levelWidth * y];
}

} // namespace

void buildMipPyramid(const HostBitmap &bmp, int filterType, int wrapMode, std::vector<float> &texels, std::vector<uint32_t> &levels) {
	const int width = bmp.width, height = bmp.height;
	int mW = width, mH = height;
	std::vector<C3> texture((size_t) width * height);
	for (size_t i = 0; i < texture.size(); ++i)          // MIPMap::fromBitmap (mipmap.cpp:169-190): clampNegative
		texture[i] = { std::max(bmp.rgb[3 * i], 0.0f), std::max(bmp.rgb[3 * i + 1], 0.0f), std::max(bmp.rgb[3 * i + 2], 0.0f) };

	if (filterType != FILTER_NONE && (!isPow2(width) || !isPow2(height))) {
		mW = (int) roundToPow2((uint32_t) width); mH = (int) roundToPow2((uint32_t) height);
		const std::vector<C3> pixels = texture;
		std::vector<C3> texture1((size_t) mW * height);
		std::vector<ResampleWeight> weights = resampleWeights(width, mW);
		for (int y = 0; y < height; ++y)
			for (int x = 0; x < mW; ++x) {
				C3 acc = { 0.0f, 0.0f, 0.0f };
				for (int j = 0; j < 4; ++j) {
					int pos = weights[x].firstTexel + j;
					if (pos < 0 || pos >= height) {                      // sic: the reference tests against the height here
						if (wrapMode == WRAP_REPEAT) pos = modulo(pos, width);
						else if (wrapMode == WRAP_CLAMP) pos = clampi(pos, 0, width - 1);
					}
					if (pos >= 0 && pos < width)
						acc = add(acc, mul(pixels[pos + (size_t) y * width], weights[x].weight[j]));
				}
				texture1[x + (size_t) mW * y] = acc;
			}
		texture.assign((size_t) mW * mH, C3{ 0.0f, 0.0f, 0.0f });
		weights = resampleWeights(height, mH);
		for (int x = 0; x < mW; ++x)
			for (int y = 0; y < mH; ++y)
				for (int j = 0; j < 4; ++j) {
					int pos = weights[y].firstTexel + j;
					if (pos < 0 || pos >= height) {
						if (wrapMode == WRAP_REPEAT) pos = modulo(pos, height);
						else if (wrapMode == WRAP_CLAMP) pos = clampi(pos, 0, height - 1);
					}
					if (pos >= 0 && pos < height) {
						C3 &dst = texture[x + (size_t) mW * y];
						dst = add(dst, mul(texture1[x + (size_t) pos * mW], weights[y].weight[j]));
					}
				}
		for (C3 &c : texture) c = { std::max(c.r, 0.0f), std::max(c.g, 0.0f), std::max(c.b, 0.0f) };
	}

	const int nLevels = filterType != FILTER_NONE ? 1 + log2i((uint32_t) std::max(width, height)) : 1;
	auto emit = [&](const std::vector<C3> &lvl, int w, int h) {
		levels.push_back((uint32_t) (texels.size() / 4)); levels.push_back((uint32_t) w); levels.push_back((uint32_t) h); levels.push_back(0u);
		for (const C3 &c : lvl) { texels.push_back(c.r); texels.push_back(c.g); texels.push_back(c.b); texels.push_back(0.0f); }
	};
	emit(texture, mW, mH);
	int lw = mW, lh = mH;
	for (int i = 1; i < nLevels; ++i) {
		const int nw = std::max(1, lw / 2), nh = std::max(1, lh / 2);
		std::vector<C3> next((size_t) nw * nh);
		for (int y = 0; y < nh; ++y)
			for (int x = 0; x < nw; ++x)
				next[x + (size_t) y * nw] = mul(add(add(add(getTexel(texture, lw, lh, wrapMode, 2 * x, 2 * y),
					getTexel(texture, lw, lh, wrapMode, 2 * x + 1, 2 * y)), getTexel(texture, lw, lh, wrapMode, 2 * x, 2 * y + 1)),
					getTexel(texture, lw, lh, wrapMode, 2 * x + 1, 2 * y + 1)), 0.25f);
		emit(next, nw, nh);
		texture.swap(next); lw = nw; lh = nh;
	}
}

void mipWeightLut(float *lut) {
	for (int i = 0; i < 128; ++i) {
		const float pos = (float) i / (float) (128 - 1);
		lut[i] = std::exp(-2.0f * pos) - std::exp(-2.0f);
	}
}

// Device functions for the non-BSDF stages of the hot path: counter-based sampler, perspective
// camera, intersection record, area-luminaire sampling. Compiled with --fmad=false (see bsdf.cuh).
#pragma once
#include "types.cuh"

// ---- sampler ----------------------------------------------------------------------------------
// Stands in for IndependentSampler/Random (src/samplers/independent.cpp:72-81,
// src/libcore/random.cpp:141-228). The reference's MT19937-64 stream depends on which worker
// thread renders which block, so it is not reproducible even between two CPU runs; what is kept is
// (a) independent uniform samples, (b) the uint->float mapping of Random::nextFloat
// (random.cpp:218-227: 23 mantissa bits, [0,1)), (c) the per-sample consumption order.
// Streams are keyed by (pixel, global sample index, seed) so the image does not depend on how
// samples are split over batches or GPUs.
__host__ __device__ inline uint64_t splitmix64(uint64_t z) {
	z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
	z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
	return z ^ (z >> 31);
}
__host__ __device__ inline float uintToUnitFloat(uint32_t u) {
	union { uint32_t u; float f; } c;
	c.u = (u >> 9) | 0x3f800000u;
	return c.f - 1.0f;
}
// Low-discrepancy mode (src/samplers/ldsampler.cpp:104-143): per pixel and draw number a scrambled (0,2)-sequence of
// sppTotal points (van der Corput in x, Sobol' dimension 2 in y), visited in a per-pixel random order. The reference
// draws the scrambles and the shuffle from its MT19937 stream; here both are functions of (pixel, draw, seed): the
// scramble words by hashing, the order by a keyed bijection of the sample index (invertible steps on log2(N) bits).
struct SamplerConfig {
	uint32_t type;        // mtsgpu_sampler_type
	uint32_t depth;       // ldsampler depth
	uint32_t log2Count;   // log2(sppTotal) in LD mode
	uint32_t lensDraw;    // 1 when the camera takes a lens sample first (renderBlock, integrator.cpp:216-224)
	uint32_t resolution;  // stratified: strata per axis (resolution^2 = sppTotal); log2Count = bits that hold sppTotal-1
	uint32_t timeDraw;    // 1 when the camera takes a time sample (Camera::needsTimeSample: shutterOpenTime > 0; integrator.cpp:159-160)
};
__host__ __device__ inline float ldVanDerCorput(uint32_t n, uint32_t scramble) {       // ldsampler.cpp:104-112
	n = (n << 16) | (n >> 16);
	n = ((n & 0x00ff00ffu) << 8) | ((n & 0xff00ff00u) >> 8);
	n = ((n & 0x0f0f0f0fu) << 4) | ((n & 0xf0f0f0f0u) >> 4);
	n = ((n & 0x33333333u) << 2) | ((n & 0xccccccccu) >> 2);
	n = ((n & 0x55555555u) << 1) | ((n & 0xaaaaaaaau) >> 1);
	n ^= scramble;
	return (float) n / 4294967296.0f;
}
__host__ __device__ inline float ldSobol2(uint32_t n, uint32_t scramble) {              // ldsampler.cpp:114-118
	for (uint32_t v = 1u << 31; n != 0; n >>= 1, v ^= v >> 1)
		if (n & 1u) scramble ^= v;
	return (float) scramble / 4294967296.0f;
}
__host__ __device__ inline uint32_t ldPermute(uint32_t i, uint32_t bits, uint32_t key) {
	if (bits == 0) return 0;
	const uint32_t mask = bits >= 32 ? 0xFFFFFFFFu : ((1u << bits) - 1u);
	const uint32_t half = (bits + 1) / 2;
	for (int round = 0; round < 4; ++round) {
		i = (i + (key >> (round * 7))) & mask;          // each line is a bijection on `bits` bits
		i = (i * (0x9E3779B1u | 1u)) & mask;
		i ^= i >> half;
		i = (i * ((key << 1) | 1u)) & mask;
		i ^= i >> half;
		key = key * 0x85EBCA6Bu + 0xC2B2AE35u;
	}
	return i;
}

#define DIM_LENS_ 0u
#define DIM_PIXEL_ 2u
// Halton (src/samplers/halton.cpp:68-84): the d-th value drawn for pixel sample i is radicalInverse(d-th prime, i);
// the first 1000 primes = the reference's primeTable (include/mitsuba/core/util.h:196, src/libcore/util.cpp:64)
#define MTS_PRIME_TABLE_SIZE 1000
__device__ __constant__ int c_primes[MTS_PRIME_TABLE_SIZE] = {
	2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37, 41, 43, 47, 53, 59, 61, 67, 71, 73, 79, 83, 89, 97,
	101, 103, 107, 109, 113, 127, 131, 137, 139, 149, 151, 157, 163, 167, 173, 179, 181, 191, 193, 197, 199, 211, 223, 227, 229,
	233, 239, 241, 251, 257, 263, 269, 271, 277, 281, 283, 293, 307, 311, 313, 317, 331, 337, 347, 349, 353, 359, 367, 373, 379,
	383, 389, 397, 401, 409, 419, 421, 431, 433, 439, 443, 449, 457, 461, 463, 467, 479, 487, 491, 499, 503, 509, 521, 523, 541,
	547, 557, 563, 569, 571, 577, 587, 593, 599, 601, 607, 613, 617, 619, 631, 641, 643, 647, 653, 659, 661, 673, 677, 683, 691,
	701, 709, 719, 727, 733, 739, 743, 751, 757, 761, 769, 773, 787, 797, 809, 811, 821, 823, 827, 829, 839, 853, 857, 859, 863,
	877, 881, 883, 887, 907, 911, 919, 929, 937, 941, 947, 953, 967, 971, 977, 983, 991, 997, 1009, 1013, 1019, 1021, 1031, 1033, 1039,
	1049, 1051, 1061, 1063, 1069, 1087, 1091, 1093, 1097, 1103, 1109, 1117, 1123, 1129, 1151, 1153, 1163, 1171, 1181, 1187, 1193, 1201, 1213, 1217, 1223,
	1229, 1231, 1237, 1249, 1259, 1277, 1279, 1283, 1289, 1291, 1297, 1301, 1303, 1307, 1319, 1321, 1327, 1361, 1367, 1373, 1381, 1399, 1409, 1423, 1427,
	1429, 1433, 1439, 1447, 1451, 1453, 1459, 1471, 1481, 1483, 1487, 1489, 1493, 1499, 1511, 1523, 1531, 1543, 1549, 1553, 1559, 1567, 1571, 1579, 1583,
	1597, 1601, 1607, 1609, 1613, 1619, 1621, 1627, 1637, 1657, 1663, 1667, 1669, 1693, 1697, 1699, 1709, 1721, 1723, 1733, 1741, 1747, 1753, 1759, 1777,
	1783, 1787, 1789, 1801, 1811, 1823, 1831, 1847, 1861, 1867, 1871, 1873, 1877, 1879, 1889, 1901, 1907, 1913, 1931, 1933, 1949, 1951, 1973, 1979, 1987,
	1993, 1997, 1999, 2003, 2011, 2017, 2027, 2029, 2039, 2053, 2063, 2069, 2081, 2083, 2087, 2089, 2099, 2111, 2113, 2129, 2131, 2137, 2141, 2143, 2153,
	2161, 2179, 2203, 2207, 2213, 2221, 2237, 2239, 2243, 2251, 2267, 2269, 2273, 2281, 2287, 2293, 2297, 2309, 2311, 2333, 2339, 2341, 2347, 2351, 2357,
	2371, 2377, 2381, 2383, 2389, 2393, 2399, 2411, 2417, 2423, 2437, 2441, 2447, 2459, 2467, 2473, 2477, 2503, 2521, 2531, 2539, 2543, 2549, 2551, 2557,
	2579, 2591, 2593, 2609, 2617, 2621, 2633, 2647, 2657, 2659, 2663, 2671, 2677, 2683, 2687, 2689, 2693, 2699, 2707, 2711, 2713, 2719, 2729, 2731, 2741,
	2749, 2753, 2767, 2777, 2789, 2791, 2797, 2801, 2803, 2819, 2833, 2837, 2843, 2851, 2857, 2861, 2879, 2887, 2897, 2903, 2909, 2917, 2927, 2939, 2953,
	2957, 2963, 2969, 2971, 2999, 3001, 3011, 3019, 3023, 3037, 3041, 3049, 3061, 3067, 3079, 3083, 3089, 3109, 3119, 3121, 3137, 3163, 3167, 3169, 3181,
	3187, 3191, 3203, 3209, 3217, 3221, 3229, 3251, 3253, 3257, 3259, 3271, 3299, 3301, 3307, 3313, 3319, 3323, 3329, 3331, 3343, 3347, 3359, 3361, 3371,
	3373, 3389, 3391, 3407, 3413, 3433, 3449, 3457, 3461, 3463, 3467, 3469, 3491, 3499, 3511, 3517, 3527, 3529, 3533, 3539, 3541, 3547, 3557, 3559, 3571,
	3581, 3583, 3593, 3607, 3613, 3617, 3623, 3631, 3637, 3643, 3659, 3671, 3673, 3677, 3691, 3697, 3701, 3709, 3719, 3727, 3733, 3739, 3761, 3767, 3769,
	3779, 3793, 3797, 3803, 3821, 3823, 3833, 3847, 3851, 3853, 3863, 3877, 3881, 3889, 3907, 3911, 3917, 3919, 3923, 3929, 3931, 3943, 3947, 3967, 3989,
	4001, 4003, 4007, 4013, 4019, 4021, 4027, 4049, 4051, 4057, 4073, 4079, 4091, 4093, 4099, 4111, 4127, 4129, 4133, 4139, 4153, 4157, 4159, 4177, 4201,
	4211, 4217, 4219, 4229, 4231, 4241, 4243, 4253, 4259, 4261, 4271, 4273, 4283, 4289, 4297, 4327, 4337, 4339, 4349, 4357, 4363, 4373, 4391, 4397, 4409,
	4421, 4423, 4441, 4447, 4451, 4457, 4463, 4481, 4483, 4493, 4507, 4513, 4517, 4519, 4523, 4547, 4549, 4561, 4567, 4583, 4591, 4597, 4603, 4621, 4637,
	4639, 4643, 4649, 4651, 4657, 4663, 4673, 4679, 4691, 4703, 4721, 4723, 4729, 4733, 4751, 4759, 4783, 4787, 4789, 4793, 4799, 4801, 4813, 4817, 4831,
	4861, 4871, 4877, 4889, 4903, 4909, 4919, 4931, 4933, 4937, 4943, 4951, 4957, 4967, 4969, 4973, 4987, 4993, 4999, 5003, 5009, 5011, 5021, 5023, 5039,
	5051, 5059, 5077, 5081, 5087, 5099, 5101, 5107, 5113, 5119, 5147, 5153, 5167, 5171, 5179, 5189, 5197, 5209, 5227, 5231, 5233, 5237, 5261, 5273, 5279,
	5281, 5297, 5303, 5309, 5323, 5333, 5347, 5351, 5381, 5387, 5393, 5399, 5407, 5413, 5417, 5419, 5431, 5437, 5441, 5443, 5449, 5471, 5477, 5479, 5483,
	5501, 5503, 5507, 5519, 5521, 5527, 5531, 5557, 5563, 5569, 5573, 5581, 5591, 5623, 5639, 5641, 5647, 5651, 5653, 5657, 5659, 5669, 5683, 5689, 5693,
	5701, 5711, 5717, 5737, 5741, 5743, 5749, 5779, 5783, 5791, 5801, 5807, 5813, 5821, 5827, 5839, 5843, 5849, 5851, 5857, 5861, 5867, 5869, 5879, 5881,
	5897, 5903, 5923, 5927, 5939, 5953, 5981, 5987, 6007, 6011, 6029, 6037, 6043, 6047, 6053, 6067, 6073, 6079, 6089, 6091, 6101, 6113, 6121, 6131, 6133,
	6143, 6151, 6163, 6173, 6197, 6199, 6203, 6211, 6217, 6221, 6229, 6247, 6257, 6263, 6269, 6271, 6277, 6287, 6299, 6301, 6311, 6317, 6323, 6329, 6337,
	6343, 6353, 6359, 6361, 6367, 6373, 6379, 6389, 6397, 6421, 6427, 6449, 6451, 6469, 6473, 6481, 6491, 6521, 6529, 6547, 6551, 6553, 6563, 6569, 6571,
	6577, 6581, 6599, 6607, 6619, 6637, 6653, 6659, 6661, 6673, 6679, 6689, 6691, 6701, 6703, 6709, 6719, 6733, 6737, 6761, 6763, 6779, 6781, 6791, 6793,
	6803, 6823, 6827, 6829, 6833, 6841, 6857, 6863, 6869, 6871, 6883, 6899, 6907, 6911, 6917, 6947, 6949, 6959, 6961, 6967, 6971, 6977, 6983, 6991, 6997,
	7001, 7013, 7019, 7027, 7039, 7043, 7057, 7069, 7079, 7103, 7109, 7121, 7127, 7129, 7151, 7159, 7177, 7187, 7193, 7207, 7211, 7213, 7219, 7229, 7237,
	7243, 7247, 7253, 7283, 7297, 7307, 7309, 7321, 7331, 7333, 7349, 7351, 7369, 7393, 7411, 7417, 7433, 7451, 7457, 7459, 7477, 7481, 7487, 7489, 7499,
	7507, 7517, 7523, 7529, 7537, 7541, 7547, 7549, 7559, 7561, 7573, 7577, 7583, 7589, 7591, 7603, 7607, 7621, 7639, 7643, 7649, 7669, 7673, 7681, 7687,
	7691, 7699, 7703, 7717, 7723, 7727, 7741, 7753, 7757, 7759, 7789, 7793, 7817, 7823, 7829, 7841, 7853, 7867, 7873, 7877, 7879, 7883, 7901, 7907, 7919 };
__device__ inline float radicalInverse(int b, uint32_t i) {                 // src/libcore/util.cpp:740-750
	const float invB = (float) 1 / (float) b;
	float x = 0.0f, f = invB;
	while (i) {
		x = __fadd_rn(x, __fmul_rn(f, (float) (i % b)));
		i /= b;
		f = __fmul_rn(f, invB);
	}
	return x;
}

struct Sampler {
	uint64_t key, pixelKey;
	uint32_t sample;
	uint32_t rouletteDraws;   // halton: 1D values this path has consumed so far (they shift every later value's prime)
	bool overflow;            // a halton / hammersley draw went past the prime table
	SamplerConfig cfg;
	__host__ __device__ void init(uint32_t pixel, uint32_t sampleIndex, uint32_t seed, SamplerConfig config) {
		key = splitmix64((((uint64_t) pixel << 32) | sampleIndex) + 0x9E3779B97F4A7C15ull * (uint64_t) (seed + 1));
		pixelKey = splitmix64((((uint64_t) pixel << 32) | 0xFFFFFFFFu) + 0xD1B54A32D192ED03ull * (uint64_t) (seed + 1));
		sample = sampleIndex; cfg = config; rouletteDraws = 0; overflow = false;
	}
	// halton: index of the first of the values a dimension consumes (renderBlock draws lens and pixel, then every
	// bounce draws luminaire 2D, BSDF 2D and — from rrDepth on, unless the bounce was a transmission — roulette 1D)
	__device__ uint32_t haltonValue(uint32_t dim) const {
		if (dim == DIM_LENS_) return 0;
		if (dim == DIM_PIXEL_) return 2 * cfg.lensDraw + cfg.timeDraw;          // lens 2D, time 1D, then the pixel sample
		const uint32_t b = (dim - 3u) / 3u, k = (dim - 3u) % 3u;
		return 2 * cfg.lensDraw + cfg.timeDraw + 2 + 4 * b + rouletteDraws + 2 * k;
	}
	// halton.cpp:74-84: "Lookup depth exceeds the prime number table size!" stops the reference's render; here the value
	// index is clamped and `overflow` set, which the kernels turn into an error flag (the call fails with MTSGPU_ERR_LIMIT)
	__device__ uint32_t clampValue(uint32_t v, uint32_t last) { if (v > last) { overflow = true; return last; } return v; }
#ifdef __CUDA_ARCH__
	// value number v of this pixel sample: halton.cpp:68-70 — radicalInverse(prime v, i); hammersley.cpp:72-79 — i / N for
	// the first value, radicalInverse(prime v-1, i) after it (cfg.resolution carries N)
	__device__ float qmcValue(uint32_t v) const {
		if (cfg.type == MTSGPU_SAMPLER_HAMMERSLEY) {
			if (v == 0) return (float) sample * (1.0f / (float) cfg.resolution);
			return radicalInverse(c_primes[v - 1], sample);
		}
		return radicalInverse(c_primes[v], sample);
	}
#endif
	// which next2D() / next1D() call of the pixel sample a dimension is (renderBlock, then Li)
	__host__ __device__ uint32_t draw2D(uint32_t dim) const {
		if (dim == DIM_LENS_) return 0;
		if (dim == DIM_PIXEL_) return cfg.lensDraw;
		return cfg.lensDraw + 1 + ((dim - 3u) / 3u) * 2u + ((dim - 3u) % 3u);        // bounce b: luminaire, then BSDF
	}
	// dimension pair `dim`: independent of call order, so no counter has to live in the path state
	__host__ __device__ void get2D(uint32_t dim, float &x, float &y) {
		const uint64_t z = splitmix64(key + 0x9E3779B97F4A7C15ull * (uint64_t) (dim + 1));
		x = uintToUnitFloat((uint32_t) z);
		y = uintToUnitFloat((uint32_t) (z >> 32));
		if (cfg.type == MTSGPU_SAMPLER_INDEPENDENT) return;
#ifdef __CUDA_ARCH__
		if (cfg.type == MTSGPU_SAMPLER_HALTON || cfg.type == MTSGPU_SAMPLER_HAMMERSLEY) {
			// next2D() is `Point2(nextValue(), nextValue())` (halton.cpp:80-84, hammersley.cpp:94-99): the order of the
			// two calls is the compiler's choice, and GCC (the reference's Linux build, and the oracle's) evaluates
			// arguments right to left — y takes the first value, x the second
			const uint32_t v = clampValue(haltonValue(dim), MTS_PRIME_TABLE_SIZE - 2u);
			y = qmcValue(v); x = qmcValue(v + 1);
			return;
		}
#endif
		const uint32_t draw = draw2D(dim);
		if (draw >= cfg.depth) return;
		if (cfg.type == MTSGPU_SAMPLER_LD) { ld2D(draw, sample, x, y); return; }
		// stratified.cpp:166-176: stratum k of a per-pixel random order, jittered
		const uint32_t k = stratum(pixelKey + 0x9E3779B97F4A7C15ull * (uint64_t) (2 * draw + 2));
		const float invRes = 1.0f / (float) cfg.resolution;
		x = ((float) (k % cfg.resolution) + x) * invRes;
		y = ((float) (k / cfg.resolution) + y) * invRes;
	}
	// this sample's entry in a keyed permutation of 0 .. resolution^2-1 (cycle-walking over the next power of two)
	__host__ __device__ uint32_t stratum(uint64_t seedKey) const {
		const uint32_t order = (uint32_t) splitmix64(seedKey), count = cfg.resolution * cfg.resolution;
		uint32_t k = sample;
		do { k = ldPermute(k, cfg.log2Count, order); } while (k >= count);
		return k;
	}
	__host__ __device__ void ld2D(uint32_t draw, uint32_t sampleIndex, float &x, float &y) const {
		const uint64_t h = splitmix64(pixelKey + 0x9E3779B97F4A7C15ull * (uint64_t) (2 * draw + 1));
		const uint64_t order = splitmix64(pixelKey + 0x9E3779B97F4A7C15ull * (uint64_t) (2 * draw + 2));
		const uint32_t n = ldPermute(sampleIndex, cfg.log2Count, (uint32_t) order);
		x = ldVanDerCorput(n, (uint32_t) h);
		y = ldSobol2(n, (uint32_t) (h >> 32));
	}
	// 1D draws: only the roulette decision uses them; draw number = how far past rrDepth the path is
	__host__ __device__ float get1D(uint32_t dim, uint32_t draw1D = 0xFFFFFFFFu) {
#ifdef __CUDA_ARCH__
		if (cfg.type == MTSGPU_SAMPLER_HALTON || cfg.type == MTSGPU_SAMPLER_HAMMERSLEY)
			return qmcValue(clampValue(haltonValue(dim), MTS_PRIME_TABLE_SIZE - 1u));
#endif
		if (draw1D != 0xFFFFFFFFu) draw1D += cfg.timeDraw;          // the time sample is the pixel sample's first 1D draw
		if (cfg.type == MTSGPU_SAMPLER_STRATIFIED && draw1D < cfg.depth) {         // stratified.cpp:157-161
			const uint64_t z = splitmix64(key + 0x9E3779B97F4A7C15ull * (uint64_t) (dim + 1));
			const uint32_t k = stratum(pixelKey + 0xA0761D6478BD642Full * (uint64_t) (2 * draw1D + 2));
			return ((float) k + uintToUnitFloat((uint32_t) z)) * (1.0f / (float) (cfg.resolution * cfg.resolution));
		}
		if (cfg.type == MTSGPU_SAMPLER_LD && draw1D < cfg.depth) {
			const uint64_t h = splitmix64(pixelKey + 0xA0761D6478BD642Full * (uint64_t) (2 * draw1D + 1));
			const uint64_t order = splitmix64(pixelKey + 0xA0761D6478BD642Full * (uint64_t) (2 * draw1D + 2));
			return ldVanDerCorput(ldPermute(sample, cfg.log2Count, (uint32_t) order), (uint32_t) h);
		}
		float x, y;
		const uint64_t z = splitmix64(key + 0x9E3779B97F4A7C15ull * (uint64_t) (dim + 1));
		x = uintToUnitFloat((uint32_t) z); y = 0; (void) y;
		return x;
	}
};
// dimension layout per pixel sample: 0 lens, 1 time, 2 pixel, then per bounce b (depth b+1):
#define DIM_LENS 0u
#define DIM_TIME 1u
#define DIM_PIXEL 2u
#define DIM_BOUNCE(b, k) (3u + 3u * (b) + (k))   // k: 0 luminaire 2D, 1 BSDF 2D, 2 roulette 1D
#define DIM_DIRECT(round, k) (3u + 3u * (round) + (k))

// ---- camera: PerspectiveCameraImpl::generateRay (src/cameras/perspective.cpp:77-112) ----------
__device__ inline void squareToDiskConcentric(float sx, float sy, float &ox, float &oy) {   // util.cpp:626-649
	const float r1 = 2.0f * sx - 1.0f, r2 = 2.0f * sy - 1.0f;
	float cx, cy;
	if (r1 == 0 && r2 == 0) { cx = 0; cy = 0; }
	else if (r1 > -r2) {
		if (r1 > r2) { cx = r1; cy = (MTS_PI / 4.0f) * r2 / r1; }
		else { cx = r2; cy = (MTS_PI / 4.0f) * (2.0f - r1 / r2); }
	} else {
		if (r1 < r2) { cx = -r1; cy = (MTS_PI / 4.0f) * (4.0f + r2 / r1); }
		else { cx = -r2; cy = (MTS_PI / 4.0f) * (6.0f - r1 / r2); }
	}
	ox = cx * mts_cos(cy); oy = cx * mts_sin(cy);
}

__device__ inline void cameraRay(const CameraDev &cam, float sx, float sy, float lx, float ly,
		V3 &o, V3 &d, float &mint, float &maxt) {
	const float *m = cam.rasterToCamera;
	// Transform::operator()(Point, Point&) with the projective divide (core/transform.h:133-149)
	float x = m[0] * sx + m[1] * sy + m[2] * 0.0f + m[3];
	float y = m[4] * sx + m[5] * sy + m[6] * 0.0f + m[7];
	float z = m[8] * sx + m[9] * sy + m[10] * 0.0f + m[11];
	const float w = m[12] * sx + m[13] * sy + m[14] * 0.0f + m[15];
	if (w != 1.0f) { const float rw = 1.0f / w; x *= rw; y *= rw; z *= rw; }
	V3 lo = v3(0, 0, 0), ld = v3(x, y, z);
	if (cam.orthographic) {
		// OrthographicCamera::generateRay (orthographic.cpp:104-119): origin on the image plane, direction +z
		lo = ld; ld = v3(0, 0, 1);
	} else if (cam.apertureRadius > 0.0f) {
		float px, py;
		squareToDiskConcentric(lx, ly, px, py);
		px *= cam.apertureRadius; py *= cam.apertureRadius;
		const float ft = cam.focusDepth / ld.z;
		const V3 itsFocal = lo + ld * ft;
		lo.x += px; lo.y += py;
		ld = itsFocal - lo;
	}
	if (cam.orthographic) {
		mint = 0; maxt = cam.farClip - cam.nearClip;
	} else {
		ld = normalize(ld);
		const float invZ = 1.0f / ld.z;
		mint = cam.nearClip * invZ;
		maxt = cam.farClip * invZ;
	}
	const float *c = cam.cameraToWorld;
	{
		float ox = c[0] * lo.x + c[1] * lo.y + c[2] * lo.z + c[3];
		float oy = c[4] * lo.x + c[5] * lo.y + c[6] * lo.z + c[7];
		float oz = c[8] * lo.x + c[9] * lo.y + c[10] * lo.z + c[11];
		const float ow = c[12] * lo.x + c[13] * lo.y + c[14] * lo.z + c[15];
		if (ow != 1.0f) { const float rw = 1.0f / ow; ox *= rw; oy *= rw; oz *= rw; }
		o = v3(ox, oy, oz);
	}
	d = v3(c[0] * ld.x + c[1] * ld.y + c[2] * ld.z,
	       c[4] * ld.x + c[5] * ld.y + c[6] * ld.z,
	       c[8] * ld.x + c[9] * ld.y + c[10] * ld.z);
}

// ---- intersection record: ShapeKDTree::fillIntersectionRecord<true> (skdtree.h:352-432) -----------
struct Isect {
	V3 p, geoN, shN, wi;
	float uvx, uvy;
	Frame3 sh;
	uint32_t shape;
	int materialId, emitterId;
};

// `tri` is the sorted triangle index the traversal reports. The hit's vertices and normals sit in one 96-byte record
// (three full sectors, one dependent load after the hit) instead of behind the chain TriAccel -> ShapeInfo -> index
// triple -> six half-used vertex sectors; ncu on C4 had k_shade waiting on exactly that chain (long-scoreboard
// stalls 6.9 per issue, 742 B of DRAM traffic per hit).
// Sphere::fillIntersectionRecord (src/shapes/sphere.cpp:133-173) from the sphere's shading record
__device__ inline void fillSphereIntersection(const DeviceScene &sc, const float4 &r0, const float4 &r1, const float4 &r2,
		const float4 &r3, const float4 &r4, const float4 &r5, V3 rayO, V3 rayD, float t, Isect &its) {
	const uint32_t shape = __float_as_uint(r0.w);
	const ShapeInfo si = sc.shapes[shape];
	const V3 center = v3(r0);
	const float radius = r1.x;
	its.p = rayO + rayD * t;                                                   // ray(its.t)
	const V3 rel = its.p - center;
	const V3 local = v3(r1.z * rel.x + r1.w * rel.y + r2.x * rel.z, r2.y * rel.x + r2.z * rel.y + r2.w * rel.z,
		r3.x * rel.x + r3.y * rel.y + r3.z * rel.z);                           // m_worldToObject(Vector)
	const float theta = mts_acos(fminf(fmaxf(local.z / radius, -1.0f), 1.0f));
	float phi = mts_atan2(local.y, local.x);
	if (phi < 0) phi += 2 * MTS_PI;
	its.uvx = phi * (0.5f * MTS_INV_PI); its.uvy = theta * MTS_INV_PI;
	const V3 du = v3(-local.y, local.x, 0.0f) * (2 * MTS_PI);
	const V3 dpdu = v3(r3.w * du.x + r4.x * du.y + r4.y * du.z, r4.z * du.x + r4.w * du.y + r5.x * du.z, r5.y * du.x + r5.z * du.y + r5.w * du.z);
	V3 n = normalize(its.p - center);
	const float zrad = sqrtf(local.x * local.x + local.y * local.y);
	if (zrad > 0) {
		const float invZRad = 1.0f / zrad, cosPhi = local.x * invZRad, sinPhi = local.y * invZRad;
		const V3 dv = v3(local.z * cosPhi, local.z * sinPhi, -mts_sin(theta) * radius) * MTS_PI;
		const V3 dpdv = v3(r3.w * dv.x + r4.x * dv.y + r4.y * dv.z, r4.z * dv.x + r4.w * dv.y + r5.x * dv.z, r5.y * dv.x + r5.z * dv.y + r5.w * dv.z);
		its.sh.s = normalize(dpdu); its.sh.t = normalize(dpdv);
	} else {
		const Frame3 f = makeFrame(n);
		its.sh.s = f.s; its.sh.t = f.t;
	}
	if (r1.y != 0.0f) n = n * -1.0f;                                           // m_inverted
	its.sh.n = n; its.geoN = n; its.shN = n;
	its.wi = toLocal(its.sh, -rayD);
	its.shape = shape; its.materialId = si.materialId; its.emitterId = si.emitterId;
}

// Triangle part of ShapeKDTree::fillIntersectionRecord<BARY> (skdtree.h:352-432) from the six float4 of a shading record
template <bool BARY>
__device__ __forceinline__ void fillMeshIntersection(const DeviceScene &sc, const ShapeInfo &si, uint32_t shape, const float4 &r0, const float4 &r1,
		const float4 &r2, const float4 &r3, const float4 &r4, const float4 &r5, float u, float v, V3 rayO, V3 rayD, float t, Isect &its);

// Instance::fillIntersectionRecord (instance.cpp:88-98): the group's own record for the nested triangle — position
// ray(t) on the WORLD ray (fillIntersectionRecord<false>), everything else in object space — then the frames taken to
// world space: normals through the inverse transpose, tangents through the matrix, each renormalised; the geometric
// frame is rebuilt around its transformed normal; wi is expressed in the transformed (not necessarily orthogonal) frame.
__device__ __noinline__ void fillInstanceIntersection(const DeviceScene &sc, const ShapeInfo &outer, uint32_t shape, uint32_t nested,
		float u, float v, V3 rayO, V3 rayD, float t, Isect &its) {
	const InstanceDev &inst = sc.instances[outer.cdfOffset];
	const DeviceScene &gs = sc.groups[inst.group];
	const float4 *rec = gs.shadeTris + (size_t) nested * 6;
	const float4 r0 = __ldg(rec), r1 = __ldg(rec + 1), r2 = __ldg(rec + 2), r3 = __ldg(rec + 3), r4 = __ldg(rec + 4), r5 = __ldg(rec + 5);
	const uint32_t gshape = __float_as_uint(r0.w);
	const ShapeInfo gsi = gs.shapes[gshape];
	fillMeshIntersection<false>(gs, gsi, gshape, r0, r1, r2, r3, r4, r5, u, v, rayO, rayD, t, its);
	const float *M = inst.o2w, *I = inst.w2o;
	auto xfVec = [&](V3 a) { return v3(M[0] * a.x + M[1] * a.y + M[2] * a.z, M[4] * a.x + M[5] * a.y + M[6] * a.z, M[8] * a.x + M[9] * a.y + M[10] * a.z); };
	auto xfNrm = [&](V3 a) { return v3(I[0] * a.x + I[4] * a.y + I[8] * a.z, I[1] * a.x + I[5] * a.y + I[9] * a.z, I[2] * a.x + I[6] * a.y + I[10] * a.z); };
	its.sh.n = normalize(xfNrm(its.sh.n));
	its.sh.s = normalize(xfVec(its.sh.s));
	its.sh.t = normalize(xfVec(its.sh.t));
	its.shN = its.sh.n;
	its.geoN = normalize(xfNrm(its.geoN));
	its.wi = toLocal(its.sh, -rayD);
	its.shape = shape; its.emitterId = -1;
}

// Cylinder::fillIntersectionRecord (src/shapes/cylinder.cpp:144-165)
__device__ __noinline__ void fillCylinderIntersection(const InstanceDev *instances, const ShapeInfo si, uint32_t shape, V3 rayO, V3 rayD, float t, Isect &its) {
	const InstanceDev &c = instances[si.cdfOffset];
	const float *M = c.o2w, *I = c.w2o;
	its.p = rayO + rayD * t;
	float lx = I[0] * its.p.x + I[1] * its.p.y + I[2] * its.p.z + I[3];
	float ly = I[4] * its.p.x + I[5] * its.p.y + I[6] * its.p.z + I[7];
	float lz = I[8] * its.p.x + I[9] * its.p.y + I[10] * its.p.z + I[11];
	const float lw = I[12] * its.p.x + I[13] * its.p.y + I[14] * its.p.z + I[15];
	if (lw != 1.0f) { const float rw = 1.0f / lw; lx *= rw; ly *= rw; lz *= rw; }
	float phi = mts_atan2(ly, lx);
	if (phi < 0) phi += 2 * MTS_PI;
	its.uvx = lz / c.length;
	its.uvy = phi / (2 * MTS_PI);
	const V3 du = v3(-ly, lx, 0.0f) * (2 * MTS_PI), dv = v3(0.0f, 0.0f, c.length);
	auto xfVec = [&](V3 a) { return v3(M[0] * a.x + M[1] * a.y + M[2] * a.z, M[4] * a.x + M[5] * a.y + M[6] * a.z, M[8] * a.x + M[9] * a.y + M[10] * a.z); };
	const V3 dpdu = xfVec(du), dpdv = xfVec(dv);
	const V3 n = normalize(xfVec(cross(du, dv)));        // Normal(normalize(objectToWorld(Vector cross(dpdu, dpdv))))
	its.geoN = n; its.shN = n;
	its.sh.n = n; its.sh.s = normalize(dpdu); its.sh.t = normalize(dpdv);
	its.wi = toLocal(its.sh, -rayD);
	its.shape = shape; its.materialId = si.materialId; its.emitterId = -1;
}

// INST: the scene may hold instances or cylinders (kernels of scenes without them are compiled without those branches)
template <bool INST = true>
__device__ inline void fillIntersectionSorted(const DeviceScene &sc, uint32_t tri, float u, float v, V3 rayO, V3 rayD, float t, Isect &its,
		uint32_t nested = 0xFFFFFFFFu) {
	const float4 *rec = sc.shadeTris + (size_t) tri * 6;
	const float4 r0 = __ldg(rec), r1 = __ldg(rec + 1), r2 = __ldg(rec + 2), r3 = __ldg(rec + 3), r4 = __ldg(rec + 4), r5 = __ldg(rec + 5);
	const uint32_t shape = __float_as_uint(r0.w);
	const ShapeInfo si = sc.shapes[shape];
	if (si.isSphere) { fillSphereIntersection(sc, r0, r1, r2, r3, r4, r5, rayO, rayD, t, its); return; }
	if constexpr (INST) {       // discarded, not merely dead, in the kernels of scenes without such shapes (a plain `if` cost them 112 bytes of stack)
		if (si.isCylinder) { fillCylinderIntersection(sc.instances, si, shape, rayO, rayD, t, its); return; }
		if (si.isInstance) { fillInstanceIntersection(sc, si, shape, nested, u, v, rayO, rayD, t, its); return; }
	}
	fillMeshIntersection<true>(sc, si, shape, r0, r1, r2, r3, r4, r5, u, v, rayO, rayD, t, its);
}

template <bool BARY>
__device__ __forceinline__ void fillMeshIntersection(const DeviceScene &sc, const ShapeInfo &si, uint32_t shape, const float4 &r0, const float4 &r1,
		const float4 &r2, const float4 &r3, const float4 &r4, const float4 &r5, float u, float v, V3 rayO, V3 rayD, float t, Isect &its) {
	const uint4 idx = make_uint4(__float_as_uint(r2.w), __float_as_uint(r3.w), __float_as_uint(r4.w), shape);
	const V3 p0 = v3(r0), p1 = v3(r1), p2 = v3(r2);
	const float bx = 1 - u - v, by = u, bz = v;
	if (BARY) its.p = p0 * bx + p1 * by + p2 * bz;      // barycentric position, NOT o + t*d
	else its.p = rayO + rayD * t;                       // ray(its.t) (skdtree.h:375-376)
	V3 fn = cross(p1 - p0, p2 - p0);
	const float len = length(fn);
	if (!isZero(fn)) fn = fn * (1.0f / len);
	its.geoN = fn;
	if (si.hasNormals) {
		const V3 n0 = v3(r3), n1 = v3(r4), n2 = v3(r5);
		its.shN = normalize(n0 * bx + n1 * by + n2 * bz);
	} else {
		its.shN = fn;
	}
	bool tangentFrame = false;
	if (si.hasNormals && si.hasTangents) {     // skdtree.h:385-398: the shading frame follows the interpolated dpdu
		const V3 d0 = v3(__ldg(sc.tangents + 2 * idx.x)), d1 = v3(__ldg(sc.tangents + 2 * idx.y)), d2 = v3(__ldg(sc.tangents + 2 * idx.z));
		const V3 dpdu = d0 * bx + d1 * by + d2 * bz;
		its.sh.n = its.shN;
		its.sh.s = normalize(dpdu - its.shN * dot(its.shN, dpdu));
		its.sh.t = cross(its.shN, its.sh.s);
		tangentFrame = true;
	}
	if (si.hasTexcoords) {                     // its.uv = t0 * b.x + t1 * b.y + t2 * b.z (skdtree.h:408-413)
		const float2 t0 = __ldg(sc.texcoords + idx.x), t1 = __ldg(sc.texcoords + idx.y), t2 = __ldg(sc.texcoords + idx.z);
		its.uvx = t0.x * bx + t1.x * by + t2.x * bz;
		its.uvy = t0.y * bx + t1.y * by + t2.y * bz;
	} else {
		its.uvx = its.uvy = 0.0f;
	}
	if (!tangentFrame) its.sh = makeFrame(its.shN);
	its.wi = toLocal(its.sh, -rayD);
	its.shape = shape; its.materialId = si.materialId; its.emitterId = si.emitterId;
}

// Same, addressed as the C ABI reports hits: (shape, primitive index within the shape)
__device__ inline void fillIntersection(const DeviceScene &sc, uint32_t shape, uint32_t prim, float u, float v,
		V3 rayO, V3 rayD, float t, Isect &its) {
	const ShapeInfo si = sc.shapes[shape];
	uint32_t nested = 0xFFFFFFFFu;
#ifdef MTSGPU_HAVE_TRACE
	if (si.isInstance) {
		// the C ABI's hit names the instance only: find the nested triangle again by intersecting the group inside
		// the one-ulp interval [t-, t] — the same arithmetic as the original query, so the same triangle answers
		// (nothing else can lie in it: t was the closest hit)
		const RayIn r = { rayO.x, rayO.y, rayO.z, 0.0f, rayD.x, rayD.y, rayD.z, INFINITY };
		float t2, u2, v2;
		if (instanceTest<false>(sc.instances, sc.groups, si.cdfOffset, r, nextafterf(t, 0.0f), t, t2, u2, v2, nested)) { u = u2; v = v2; }
		if (nested == 0xFFFFFFFFu) {             // cannot happen for a hit this library reported
			its.p = its.geoN = its.shN = its.wi = v3(0, 0, 0); its.uvx = its.uvy = 0; its.sh = makeFrame(v3(0, 0, 1));
			its.shape = shape; its.materialId = -1; its.emitterId = -1;
			return;
		}
	}
#endif
	fillIntersectionSorted(sc, __ldg(sc.sortedOf + si.triOffset + ((si.isSphere || si.isInstance || si.isCylinder) ? 0u : prim)), u, v, rayO, rayD, t, its, nested);
}

// ---- textures: Texture2D::getValue (texture.cpp:73-82) + Checkerboard::getValue (checkerboard.cpp:47-55) ----------
__device__ inline int mtsModulo(int a, int b) { const int r = a % b; return (r < 0) ? r + b : r; }   // util.h modulo()
__device__ inline V3 textureValue(const TextureDev &t, float u, float v) {
	const float tu = u * t.p[8] + t.p[6], tv = v * t.p[9] + t.p[7];
	if (t.type == MTSGPU_TEXTURE_GRID) {                    // GridTexture::getValue (gridtexture.cpp:49-60); .5 is a double there
		float gx = tu - (int) tu, gy = tv - (int) tv;
		if (gx > .5f) gx -= 1;
		if (gy > .5f) gy -= 1;
		return (fabsf(gx) < t.p[10] || fabsf(gy) < t.p[10]) ? v3(t.p[3], t.p[4], t.p[5]) : v3(t.p[0], t.p[1], t.p[2]);
	}
	const int x = 2 * mtsModulo((int) (tu * 2), 2) - 1, y = 2 * mtsModulo((int) (tv * 2), 2) - 1;
	return (x * y == 1) ? v3(t.p[0], t.p[1], t.p[2]) : v3(t.p[3], t.p[4], t.p[5]);
}

// ---- bitmap textures: MIPMap (src/librender/mipmap.cpp) over the pyramid built by mipmap.cpp ---------------------
// The three arrays a lookup reads, by value: a __noinline__ callee that took the whole DeviceScene by reference would
// force the kernel's parameter block into local memory in every kernel that can reach it.
struct MipRef { const float4 *texels; const uint4 *mipLevels; const float *mipLut; };
__device__ __forceinline__ MipRef mipRef(const DeviceScene &sc) { MipRef m = { sc.texels, sc.mipLevels, sc.mipLut }; return m; }

// MIPMap::getTexel (mipmap.cpp:220-243), including its `x <= 0` bounds test
__device__ inline V3 mipTexel(const MipRef sc, const TextureDev &t, int level, int x, int y) {
	const uint4 L = __ldg(sc.mipLevels + t.pad[0] + level);
	const int w = (int) L.y, h = (int) L.z, wrap = (int) t.p[1];
	if (x <= 0 || y < 0 || x >= w || y >= h) {
		if (wrap == 0) { x = mtsModulo(x, w); y = mtsModulo(y, h); }
		else if (wrap == 3) { x = min(max(x, 0), w - 1); y = min(max(y, 0), h - 1); }
		else return wrap == 1 ? v3(0, 0, 0) : v3(1, 1, 1);
	}
	return v3(__ldg(sc.texels + L.x + (uint32_t) (x + w * y)));
}

// MIPMap::triangle (mipmap.cpp:245-262)
__device__ inline V3 mipTriangle(const MipRef sc, const TextureDev &t, int level, float x, float y) {
	if (t.p[0] == 2.0f) {
		const uint4 L = __ldg(sc.mipLevels + t.pad[0]);
		return mipTexel(sc, t, 0, (int) floorf(x * (float) (int) L.y), (int) floorf(y * (float) (int) L.z));
	}
	level = min(max(level, 0), (int) t.pad[1] - 1);
	const uint4 L = __ldg(sc.mipLevels + t.pad[0] + level);
	x = x * (float) (int) L.y - 0.5f;
	y = y * (float) (int) L.z - 0.5f;
	const int xPos = (int) floorf(x), yPos = (int) floorf(y);
	const float dx = x - xPos, dy = y - yPos;
	return mipTexel(sc, t, level, xPos, yPos) * (1.0f - dx) * (1.0f - dy)
		+ mipTexel(sc, t, level, xPos, yPos + 1) * (1.0f - dx) * dy
		+ mipTexel(sc, t, level, xPos + 1, yPos) * dx * (1.0f - dy)
		+ mipTexel(sc, t, level, xPos + 1, yPos + 1) * dx * dy;
}

// MIPMap::EWA (mipmap.cpp:318-357)
__device__ inline V3 mipEWA(const MipRef sc, const TextureDev &t, float u, float v, float dudx, float dudy, float dvdx, float dvdy, int level) {
	const int levels = (int) t.pad[1];
	if (level >= levels) return mipTexel(sc, t, levels - 1, 0, 0);
	const uint4 L = __ldg(sc.mipLevels + t.pad[0] + level);
	const float lw = (float) (int) L.y, lh = (float) (int) L.z;
	V3 result = v3(0, 0, 0);
	float denominator = 0.0f;
	u = u * lw; v = v * lh;
	dudx = dudx * lw; dudy = dudy * lh;
	dvdx = dvdx * lw; dvdy = dvdy * lh;
	float A = dudy * dudy + dvdy * dvdy + 1.0f;
	float B = -2.0f * (dudx * dudy + dvdx * dvdy);
	float C = dudx * dudx + dvdx * dvdx + 1.0f;
	const float F = A * C - B * B * 0.25f;
	const float du = sqrtf(C), dv = sqrtf(A);
	const int u0 = (int) ceilf(u - du), u1 = (int) floorf(u + du);
	const int v0 = (int) ceilf(v - dv), v1 = (int) floorf(v + dv);
	const float invF = 1.0f / F;
	A *= invF; B *= invF; C *= invF;
	for (int ut = u0; ut <= u1; ++ut) {
		const float uu = ut - u;
		for (int vt = v0; vt <= v1; ++vt) {
			const float vv = vt - v;
			const float r2 = A * uu * uu + B * uu * vv + C * vv * vv;
			if (r2 < 1) {
				const float weight = __ldg(sc.mipLut + max(0, min((int) (r2 * 128), 128 - 1)));
				result = result + mipTexel(sc, t, level, ut, vt) * weight;
				denominator += weight;
			}
		}
	}
	return result * (1.0f / denominator);
}

__device__ inline float mtsLog2(float x) { return mts_log(x) * (1.0f / 0.693147182464599609375f); }   // util.cpp:365-368, logf(2.0f)

// MIPMap::getValue (mipmap.cpp:264-316) at the uv of Texture2D::getValue (texture.cpp:73-82). duv = {dudx, dudy, dvdx, dvdy}
// of the hit; all zero for a ray without differentials, which every filter answers with the bilinear level-0 lookup.
__device__ __noinline__ V3 bitmapValue(const MipRef sc, const TextureDev t, float u, float v, float4 duv) {
	const float tu = u * t.p[8] + t.p[6], tv = v * t.p[9] + t.p[7];
	float dudx = duv.x * t.p[8], dudy = duv.y * t.p[8], dvdx = duv.z * t.p[9], dvdy = duv.w * t.p[9];
	const int levels = (int) t.pad[1];
	if (t.p[0] == 1.0f) {                                  // trilinear
		const float width = 2.0f * fmaxf(fmaxf(fabsf(dudx), fabsf(dudy)), fmaxf(fabsf(dvdx), fabsf(dvdy)));
		const float mipmapLevel = levels - 1 + mtsLog2(fmaxf(width, 1e-8f));
		if (mipmapLevel < 0) return mipTriangle(sc, t, 0, tu, tv);
		else if (mipmapLevel >= levels - 1) return mipTexel(sc, t, levels - 1, 0, 0);
		const int level = (int) mipmapLevel;
		const float delta = mipmapLevel - level;
		const V3 a = mipTriangle(sc, t, level, tu, tv);    // sic: both terms of the blend read the same level
		return a * (1.0f - delta) + a * delta;
	} else if (t.p[0] == 0.0f) {                           // EWA
		if (dudx * dudx + dudy * dudy < dvdx * dvdx + dvdy * dvdy) {
			float x = dudx; dudx = dvdx; dvdx = x;
			x = dudy; dudy = dvdy; dvdy = x;
		}
		const float majorLength = sqrtf(dudx * dudx + dudy * dudy);
		float minorLength = sqrtf(dvdx * dvdx + dvdy * dvdy);
		const float maxAnisotropy = t.p[2];
		if (minorLength * maxAnisotropy < majorLength && minorLength > 0.0f) {
			const float scale = majorLength / (minorLength * maxAnisotropy);
			dvdx *= scale; dvdy *= scale;
			minorLength *= scale;
		}
		if (minorLength == 0) return mipTriangle(sc, t, 0, tu, tv);
		const float lod = fminf(fmaxf(0.0f, levels - 1 + mtsLog2(minorLength)), (float) (levels - 1));
		const int ilod = (int) floorf(lod);
		const float d = lod - ilod;
		return mipEWA(sc, t, tu, tv, dudx, dudy, dvdx, dvdy, ilod) * (1 - d) + mipEWA(sc, t, tu, tv, dudx, dudy, dvdx, dvdy, ilod + 1) * d;
	}
	const uint4 L = __ldg(sc.mipLevels + t.pad[0]);
	return mipTexel(sc, t, 0, (int) floorf(tu * (float) (int) L.y), (int) floorf(tv * (float) (int) L.z));
}

// uv partials of a camera-ray hit. The reference carries the two offset rays of Camera::generateRayDifferential
// (src/librender/camera.cpp:79-90: the same lens sample, the raster position moved by one pixel in x / in y) along with
// the camera ray (narrowed by 1 / sqrt(spp)) and, when the hit's BSDF reads a texture that filters, solves for du/dx .. dv/dy in
// Intersection::computePartials (src/librender/intersection.cpp:5-75). Here the offset rays are rebuilt from the
// sample's raster position and lens sample, and dpdu / dpdv from the hit's shading record: this runs once per path,
// at the first hit, and only on surfaces with a bitmap texture. `tri` is the sorted triangle index of the hit.
__device__ __noinline__ float4 cameraHitPartials(const DeviceScene &sc, float sx, float sy, float lx, float ly,
		V3 rayO, V3 rayD, float diffScale, uint32_t tri, float bu, float bv, V3 p, V3 geoN, uint32_t nested = 0xFFFFFFFFu) {
	const float4 zero = make_float4(0, 0, 0, 0);
	// ---- its.dpdu, its.dpdv (skdtree.h:385-406 for meshes, sphere.cpp:146-165 for spheres)
	const float4 *rec = sc.shadeTris + (size_t) tri * 6;
	const float4 r0 = __ldg(rec), r1 = __ldg(rec + 1), r2 = __ldg(rec + 2), r3 = __ldg(rec + 3), r4 = __ldg(rec + 4), r5 = __ldg(rec + 5);
	const ShapeInfo si = sc.shapes[__float_as_uint(r0.w)];
	V3 dpdu = v3(0, 0, 0), dpdv = v3(0, 0, 0);
	if (si.isSphere) {
		const V3 rel = p - v3(r0);
		const float radius = r1.x;
		const V3 local = v3(r1.z * rel.x + r1.w * rel.y + r2.x * rel.z, r2.y * rel.x + r2.z * rel.y + r2.w * rel.z,
			r3.x * rel.x + r3.y * rel.y + r3.z * rel.z);
		const float theta = mts_acos(fminf(fmaxf(local.z / radius, -1.0f), 1.0f));
		const V3 du = v3(-local.y, local.x, 0.0f) * (2 * MTS_PI);
		dpdu = v3(r3.w * du.x + r4.x * du.y + r4.y * du.z, r4.z * du.x + r4.w * du.y + r5.x * du.z, r5.y * du.x + r5.z * du.y + r5.w * du.z);
		const float zrad = sqrtf(local.x * local.x + local.y * local.y);
		float cosPhi = 0, sinPhi = 1;
		if (zrad > 0) { const float invZRad = 1.0f / zrad; cosPhi = local.x * invZRad; sinPhi = local.y * invZRad; }
		const V3 dv = v3(local.z * cosPhi, local.z * sinPhi, -mts_sin(theta) * radius) * MTS_PI;
		dpdv = v3(r3.w * dv.x + r4.x * dv.y + r4.y * dv.z, r4.z * dv.x + r4.w * dv.y + r5.x * dv.z, r5.y * dv.x + r5.z * dv.y + r5.w * dv.z);
	} else if (si.isCylinder) {
		const InstanceDev &c = sc.instances[si.cdfOffset];
		const float *M = c.o2w, *I = c.w2o;
		const float lx = I[0] * p.x + I[1] * p.y + I[2] * p.z + I[3], ly = I[4] * p.x + I[5] * p.y + I[6] * p.z + I[7];
		const V3 du = v3(-ly, lx, 0.0f) * (2 * MTS_PI), dv = v3(0.0f, 0.0f, c.length);
		dpdu = v3(M[0] * du.x + M[1] * du.y + M[2] * du.z, M[4] * du.x + M[5] * du.y + M[6] * du.z, M[8] * du.x + M[9] * du.y + M[10] * du.z);
		dpdv = v3(M[0] * dv.x + M[1] * dv.y + M[2] * dv.z, M[4] * dv.x + M[5] * dv.y + M[6] * dv.z, M[8] * dv.x + M[9] * dv.y + M[10] * dv.z);
	} else if (si.isInstance) {
		// Instance::fillIntersectionRecord (instance.cpp:96-97): the group's dpdu / dpdv through objectToWorld
		const InstanceDev &inst = sc.instances[si.cdfOffset];
		const DeviceScene &gs = sc.groups[inst.group];
		const float4 *grec = gs.shadeTris + (size_t) nested * 6;
		const ShapeInfo gsi = gs.shapes[__float_as_uint(__ldg(grec).w)];
		if (nested != 0xFFFFFFFFu && gsi.hasNormals && gsi.hasTangents) {
			const uint32_t i0 = __float_as_uint(__ldg(grec + 2).w), i1 = __float_as_uint(__ldg(grec + 3).w), i2 = __float_as_uint(__ldg(grec + 4).w);
			const float bx = 1 - bu - bv, by = bu, bz = bv;
			const V3 du = v3(__ldg(gs.tangents + 2 * i0)) * bx + v3(__ldg(gs.tangents + 2 * i1)) * by + v3(__ldg(gs.tangents + 2 * i2)) * bz;
			const V3 dv = v3(__ldg(gs.tangents + 2 * i0 + 1)) * bx + v3(__ldg(gs.tangents + 2 * i1 + 1)) * by + v3(__ldg(gs.tangents + 2 * i2 + 1)) * bz;
			const float *M = inst.o2w;
			dpdu = v3(M[0] * du.x + M[1] * du.y + M[2] * du.z, M[4] * du.x + M[5] * du.y + M[6] * du.z, M[8] * du.x + M[9] * du.y + M[10] * du.z);
			dpdv = v3(M[0] * dv.x + M[1] * dv.y + M[2] * dv.z, M[4] * dv.x + M[5] * dv.y + M[6] * dv.z, M[8] * dv.x + M[9] * dv.y + M[10] * dv.z);
		}
	} else if (si.hasNormals && si.hasTangents) {
		const uint32_t i0 = __float_as_uint(r2.w), i1 = __float_as_uint(r3.w), i2 = __float_as_uint(r4.w);
		const float bx = 1 - bu - bv, by = bu, bz = bv;
		dpdu = v3(__ldg(sc.tangents + 2 * i0)) * bx + v3(__ldg(sc.tangents + 2 * i1)) * by + v3(__ldg(sc.tangents + 2 * i2)) * bz;
		dpdv = v3(__ldg(sc.tangents + 2 * i0 + 1)) * bx + v3(__ldg(sc.tangents + 2 * i1 + 1)) * by + v3(__ldg(sc.tangents + 2 * i2 + 1)) * bz;
	}
	if (isZero(dpdu) && isZero(dpdv)) return zero;
	// ---- the offset rays
	const CameraDev cam = *sc.camera;
	V3 rxO, rxD, ryO, ryD; float mint, maxt;
	cameraRay(cam, sx + 1, sy, lx, ly, rxO, rxD, mint, maxt);
	cameraRay(cam, sx, sy + 1, lx, ly, ryO, ryD, mint, maxt);
	// RayDifferential::scaleDifferential (core/ray.h:150-155) by 1 / sqrt(samples per pixel) (integrator.cpp:145,165)
	rxO = rayO + (rxO - rayO) * diffScale; ryO = rayO + (ryO - rayO) * diffScale;
	rxD = rayD + (rxD - rayD) * diffScale; ryD = rayD + (ryD - rayD) * diffScale;
	// ---- Intersection::computePartials
	const float d = -dot(geoN, p);
	const float txRecip = dot(geoN, rxD), tyRecip = dot(geoN, ryD);
	if (txRecip == 0 || tyRecip == 0) return zero;
	const float tx = -(dot(geoN, rxO) + d) / txRecip;
	const float ty = -(dot(geoN, ryO) + d) / tyRecip;
	const float absX = fabsf(geoN.x), absY = fabsf(geoN.y), absZ = fabsf(geoN.z);
	int a0, a1;
	if (absX > absY && absX > absZ) { a0 = 1; a1 = 2; }
	else if (absY > absZ) { a0 = 0; a1 = 2; }
	else { a0 = 0; a1 = 1; }
	auto comp = [](V3 v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : v.z); };
	const float A00 = comp(dpdu, a0), A01 = comp(dpdv, a0), A10 = comp(dpdu, a1), A11 = comp(dpdv, a1);
	const V3 px = rxO + rxD * tx, py = ryO + ryD * ty;
	const float Bx0 = comp(px, a0) - comp(p, a0), Bx1 = comp(px, a1) - comp(p, a1);
	const float By0 = comp(py, a0) - comp(p, a0), By1 = comp(py, a1) - comp(p, a1);
	// solveLinearSystem2x2 (util.cpp:490-502)
	const float det = A00 * A11 - A01 * A10;
	float4 out;
	if (fabsf(det) < MTS_EPSILON) {
		// the fall-backs of intersection.cpp:64-74; its second one assigns dudy twice and leaves dvdy unset
		// (left at 0 here)
		out = make_float4(1, 1, 0, 0);
	} else {
		const float inverse = 1.0f / det;
		out.x = (A11 * Bx0 - A01 * Bx1) * inverse;       // dudx
		out.z = (A00 * Bx1 - A10 * Bx0) * inverse;       // dvdx
		out.y = (A11 * By0 - A01 * By1) * inverse;       // dudy
		out.w = (A00 * By1 - A10 * By0) * inverse;       // dvdy
	}
	return out;
}

__device__ inline bool usesRayDifferentials(const DeviceScene &sc, const MaterialDev &m) {
	return (m.tex[0] && sc.textures[m.tex[0] - 1].type == MTSGPU_TEXTURE_BITMAP) || (m.tex[1] && sc.textures[m.tex[1] - 1].type == MTSGPU_TEXTURE_BITMAP)
		|| (m.tex[2] && sc.textures[m.tex[2] - 1].type == MTSGPU_TEXTURE_BITMAP);
}

// material with its textured slots replaced by their values at (u, v)
template <bool BMP = true>
__device__ inline void resolveTextures(const DeviceScene &sc, MaterialDev &m, float u, float v, float4 duv = make_float4(0, 0, 0, 0)) {
#pragma unroll
	for (int k = 0; k < 3; ++k)
		if (m.tex[k]) {
			const TextureDev &t = sc.textures[m.tex[k] - 1];
			const V3 c = (BMP && t.type == MTSGPU_TEXTURE_BITMAP) ? bitmapValue(mipRef(sc), t, u, v, duv) : textureValue(t, u, v);
			const int at = k == 2 ? 13 : 3 * k;              // slot 2: the opacity of a mask wrapper
			m.p[at] = c.x; m.p[at + 1] = c.y; m.p[at + 2] = c.z;
		}
}

// ---- DiscretePDF::sampleReuse (include/mitsuba/core/pdf.h:107-148) on a device CDF ----------------
// std::lower_bound: first entry >= value; index = clamp(pos - 1, 0, n - 1)
__device__ inline int cdfSampleReuse(const float *cdf, int n, float &value) {
	int lo = 0, count = n + 1;
	while (count > 0) {
		const int step = count / 2, mid = lo + step;
		if (cdf[mid] < value) { lo = mid + 1; count -= step + 1; } else count = step;
	}
	int index = max(0, lo - 1);
	index = min(index, n - 1);
	value = (value - cdf[index]) / (cdf[index + 1] - cdf[index]);
	return index;
}

// ---- Scene::sampleLuminaire without the visibility test (scene.cpp:396-415) -----------------------
struct LumSample {
	V3 p, n, d, value;   // sRec.p, sRec.n, lRec.d (normalised, light -> point), lRec.value (already / pdf)
	float pdf;           // lRec.pdf after multiplication by the luminaire-selection pdf; 0 => failed
	int emitter;
	bool mis;            // luminaire->isIntersectable() || isBackgroundLuminaire(): BSDF sampling can find it too
};

__device__ inline V3 squareToSphere(float sx, float sy) {                   // util.cpp:552-558
	const float z = 1.0f - 2.0f * sy;
	float r = 1.0f - z * z;
	r = sqrtf(fmaxf(0.0f, r));
	const float phi = 2.0f * MTS_PI * sx;
	return v3(r * mts_cos(phi), r * mts_sin(phi), z);
}

// BSphere::rayIntersect (core/bsphere.h:85-118) for a ray with mint = 0
__device__ inline bool bsphereRayIntersect(V3 center, float radius, V3 o, V3 d, float &nearHit, float &farHit) {
	const V3 originToCenter = center - o;
	const float distToRayClosest = dot(originToCenter, d);
	const float tmp1 = dot(originToCenter, originToCenter) - radius * radius;
	if (tmp1 <= 0.0f) {
		nearHit = farHit = sqrtf(distToRayClosest * distToRayClosest - tmp1) + distToRayClosest;
		return true;
	}
	if (distToRayClosest < 0.0f) return false;
	const float sqrHalfChordDist = radius * radius - dot(originToCenter, originToCenter) + distToRayClosest * distToRayClosest;
	if (sqrHalfChordDist < 0) return false;
	const float hitDistance = sqrtf(sqrHalfChordDist);
	nearHit = distToRayClosest - hitDistance; farHit = distToRayClosest + hitDistance;
	if (nearHit == 0) nearHit = farHit;
	return true;
}

// squareToCone / squareToConePdf (util.cpp:652-662)
__device__ inline V3 squareToCone(float cosCutoff, float sx, float sy) {
	const float cosTheta = (1 - sx) + sx * cosCutoff;
	const float sinTheta = sqrtf(1 - cosTheta * cosTheta);
	const float phi = sy * (2 * MTS_PI);
	return v3(mts_cos(phi) * sinTheta, mts_sin(phi) * sinTheta, cosTheta);
}

// Sphere::rayIntersect(ray, 0, inf, t) (sphere.cpp:92-112) with solveQuadratic (util.cpp:450-488)
__device__ inline bool sphereRayIntersect(V3 center, float radius, V3 ro, V3 rd, float &t) {
	const V3 o = ro - center;
	const float A = rd.x * rd.x + rd.y * rd.y + rd.z * rd.z;
	const float B = 2 * (rd.x * o.x + rd.y * o.y + rd.z * o.z);
	const float C = o.x * o.x + o.y * o.y + o.z * o.z - radius * radius;
	float nearT, farT;
	if (A == 0) {
		if (B == 0) return false;
		nearT = farT = -C / B;
	} else {
		const float discrim = B * B - 4.0f * A * C;
		if (discrim < 0) return false;
		const float sqrtDiscrim = sqrtf(discrim);
		const float temp = (B < 0) ? -0.5f * (B - sqrtDiscrim) : -0.5f * (B + sqrtDiscrim);
		nearT = temp / A; farT = C / temp;
		if (nearT > farT) { const float x = nearT; nearT = farT; farT = x; }
	}
	if (farT < 0.0f) return false;              // nearT > inf cannot happen
	t = (nearT < 0.0f) ? farT : nearT;           // farT > inf cannot happen
	return true;
}

// Sphere::sampleSolidAngle (sphere.cpp:191-229): cone sampling from outside (Shirley, Wang & Zimmerman 1996),
// uniform area sampling from inside. Returns the pdf wrt. solid angle, 0 on failure.
__device__ inline float sphereSampleSolidAngle(V3 center, float radius, float invSurfaceArea, V3 p, float sx, float sy, V3 &sp, V3 &sn) {
	const V3 w = center - p; const float invDistW = 1 / length(w);
	const float squareTerm = fabsf(radius * invDistW);
	if (squareTerm >= 1 - MTS_EPSILON) {
		const V3 d = squareToSphere(sx, sy);
		sp = center + d * radius; sn = d;
		const V3 lumToPoint = p - sp;
		const float distSquared = dot(lumToPoint, lumToPoint), dp = dot(lumToPoint, sn);
		return dp > 0 ? invSurfaceArea * distSquared * sqrtf(distSquared) / dp : 0.0f;
	}
	const float cosThetaMax = sqrtf(fmaxf(0.0f, 1 - squareTerm * squareTerm));
	const V3 d = toWorld(makeFrame(w * invDistW), squareToCone(cosThetaMax, sx, sy));
	float t;
	if (!sphereRayIntersect(center, radius, p, d, t)) return 0.0f;
	sp = p + d * t;
	sn = normalize(sp - center);
	return 1 / ((2 * MTS_PI) * (1 - cosThetaMax));
}

// Sphere::pdfSolidAngle (sphere.cpp:231-247)
__device__ inline float spherePdfSolidAngle(V3 center, float radius, float invSurfaceArea, V3 p, V3 sp, V3 sn) {
	const V3 w = p - center; const float invDistW = 1 / length(w);
	const float squareTerm = fabsf(radius * invDistW);
	if (squareTerm >= 1 - MTS_EPSILON) {
		const V3 lumToPoint = p - sp;
		const float distSquared = dot(lumToPoint, lumToPoint), dp = dot(lumToPoint, sn);
		return dp > 0 ? invSurfaceArea * distSquared * sqrtf(distSquared) / dp : 0.0f;
	}
	const float cosThetaMax = sqrtf(fmaxf(0.0f, 1 - squareTerm * squareTerm));
	return 1 / (2 * MTS_PI * (1 - cosThetaMax));
}

// ---- environment map (src/luminaires/envmap.cpp). sc.envData: worldToLuminaire 3x3, luminaireToWorld 3x3, the
// resolution of the sampling density (a coarse MIP level), its CDF (n + 1) and its normalised pdf (n).
__device__ inline V3 envXf(const float *m, V3 v) { return v3(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[3] * v.x + m[4] * v.y + m[5] * v.z, m[6] * v.x + m[7] * v.y + m[8] * v.z); }

// EnvMapLuminaire::Le(direction) (envmap.cpp:158-164): bilinear level-0 lookup at the direction's latitude / longitude
__device__ __noinline__ V3 envLe(const MipRef mr, const TextureDev tex, const float *envData, float scale, V3 direction) {
	const V3 d = envXf(envData, direction);
	const float u = .5f * (1 + mts_atan2(d.x, -d.z) / MTS_PI);
	const float v = mts_acos(fmaxf(-1.0f, fminf(1.0f, d.y))) / MTS_PI;
	return mipTriangle(mr, tex, 0, u, v) * scale;
}

// EnvMapLuminaire::pdf (envmap.cpp:184-199); `dir` is -lRec.d, the direction from the shaded point to the map
__device__ __noinline__ float envPdf(const float *envData, V3 dir) {
	const V3 d = envXf(envData, dir);
	const int resX = (int) envData[18], resY = (int) envData[19];
	const float x = .5f * (1 + mts_atan2(d.x, -d.z) / MTS_PI) * resX;
	const float y = mts_acos(fmaxf(-1.0f, fminf(1.0f, d.y))) / MTS_PI * resY;
	const int xPos = min(max((int) floorf(x), 0), resX - 1), yPos = min(max((int) floorf(y), 0), resY - 1);
	const float pdf = envData[20 + resX * resY + 1 + xPos + yPos * resX];
	const float sinTheta = sqrtf(fmaxf(MTS_EPSILON, 1 - d.y * d.y));
	const float pixX = 2 * MTS_PI / resX, pixY = MTS_PI / resY;
	return pdf / (pixX * pixY * sinTheta);
}

// EnvMapLuminaire::sampleDirection (envmap.cpp:134-156): a texel of the density, then a point inside it.
// Returns {direction, pdf}; `value` through the second result.
struct EnvSample { V3 d, value; float pdf; };
__device__ __noinline__ EnvSample envSampleDirection(const MipRef mr, const TextureDev tex, const float *envData, float scale, float sx, float sy) {
	EnvSample out;
	const int resX = (int) envData[18], resY = (int) envData[19];
	const float *cdf = envData + 20;
	const int idx = cdfSampleReuse(cdf, resX * resY, sx);
	float pdf = cdf[resX * resY + 1 + idx];
	const int row = idx / resX, col = idx - resX * row;
	const float x = col + sx, y = row + sy;
	out.value = mipTriangle(mr, tex, 0, x * (1.0f / resX), y * (1.0f / resY)) * scale;
	const float pixX = 2 * MTS_PI / resX, pixY = MTS_PI / resY;
	const float theta = pixY * y, phi = pixX * x - MTS_PI;
	const float sinTheta = mts_sin(theta), cosTheta = mts_cos(theta);
	const float sinPhi = mts_sin(phi), cosPhi = mts_cos(phi);
	out.pdf = pdf / (pixX * pixY * sinTheta);
	out.d = envXf(envData + 9, v3(-sinTheta * sinPhi, -cosTheta, sinTheta * cosPhi));
	return out;
}

// CollimatedBeamLuminaire::sample (collimated.cpp:63-75): the point in the beam's frame, the disc test, the projection
// back. `xf`: luminaireToWorld 4x4, worldToLuminaire 4x4. Out of line and by value, so that scenes without a beam pay
// nothing for it in the shading kernel.
struct BeamSample { V3 p, d; bool ok; };
__device__ __noinline__ BeamSample collimatedSample(const float *xf, float radius, V3 p) {
	const float *L = xf, *W = xf + 16;
	BeamSample out; out.ok = false; out.p = out.d = v3(0, 0, 0);
	float lx = W[0] * p.x + W[1] * p.y + W[2] * p.z + W[3], ly = W[4] * p.x + W[5] * p.y + W[6] * p.z + W[7],
		lz = W[8] * p.x + W[9] * p.y + W[10] * p.z + W[11];
	const float lw = W[12] * p.x + W[13] * p.y + W[14] * p.z + W[15];
	if (lw != 1.0f) { const float rw = 1.0f / lw; lx *= rw; ly *= rw; lz *= rw; }
	if (sqrtf(lx * lx + ly * ly) > radius || lz < 0) return out;
	float wx = L[0] * lx + L[1] * ly + L[2] * 0.0f + L[3], wy = L[4] * lx + L[5] * ly + L[6] * 0.0f + L[7],
		wz = L[8] * lx + L[9] * ly + L[10] * 0.0f + L[11];
	const float ww = L[12] * lx + L[13] * ly + L[14] * 0.0f + L[15];
	if (ww != 1.0f) { const float rw = 1.0f / ww; wx *= rw; wy *= rw; wz *= rw; }
	out.p = v3(wx, wy, wz);
	out.d = v3(L[0] * 0.0f + L[1] * 0.0f + L[2], L[4] * 0.0f + L[5] * 0.0f + L[6], L[8] * 0.0f + L[9] * 0.0f + L[10]);     // m_direction = luminaireToWorld(Vector(0, 0, 1))
	out.ok = true;
	return out;
}

__device__ inline void sampleLuminaire(const DeviceScene &sc, V3 p, float sx, float sy, bool importance, LumSample &out) {
	out.pdf = 0; out.emitter = -1; out.mis = true;
	if (sc.nEmitters == 0) return;
	const int index = cdfSampleReuse(sc.lumCdf, (int) sc.nEmitters, sx);
	const EmitterDev em = sc.emitters[index];
	const float lumPdf = em.pdf;
	const V3 inten = v3(em.intensity[0], em.intensity[1], em.intensity[2]);
	out.emitter = index;
	float pdf = 0.0f;
	V3 value = inten;
	if (em.type == MTSGPU_EMITTER_AREA) {
		const ShapeInfo si = sc.shapes[em.shape];
		if (si.isSphere) {
			const float4 cr = __ldg(sc.positions + si.vertOffset);            // centre, radius
			pdf = sphereSampleSolidAngle(v3(cr), cr.w, si.invSurfaceArea, p, sx, sy, out.p, out.n);
		} else {
		// TriMesh::sampleArea (trimesh.cpp:297-302) + Triangle::sample (triangle.cpp:23-47)
		const int tri = cdfSampleReuse(sc.areaCdf + si.cdfOffset, (int) si.nTri, sy);
		const uint4 idx = __ldg(sc.triIdx + si.triOffset + tri);
		const V3 p0 = v3(__ldg(sc.positions + idx.x)), p1 = v3(__ldg(sc.positions + idx.y)), p2 = v3(__ldg(sc.positions + idx.z));
		const float a = sqrtf(1.0f - sx);
		const float baryX = 1 - a, baryY = a * sy;
		const V3 sideA = p1 - p0, sideB = p2 - p0;
		out.p = p0 + (sideA * baryX) + (sideB * baryY);
		if (si.hasNormals) {
			const V3 n0 = v3(__ldg(sc.normals + idx.x)), n1 = v3(__ldg(sc.normals + idx.y)), n2 = v3(__ldg(sc.normals + idx.z));
			out.n = normalize(n0 * (1.0f - baryX - baryY) + n1 * baryX + n2 * baryY);
		} else {
			out.n = normalize(cross(sideA, sideB));
		}
		// Shape::sampleSolidAngle (shape.cpp:65-75)
		const V3 lumToPoint = p - out.p;
		const float distSquared = dot(lumToPoint, lumToPoint), dp = dot(lumToPoint, out.n);
		if (dp > 0)
			pdf = si.invSurfaceArea * distSquared * sqrtf(distSquared) / dp;
		}
		// AreaLuminaire::sample (area.cpp:68-79)
		V3 d = p - out.p;
		if (pdf > 0 && dot(d, out.n) > 0) {
			d = normalize(d);
		} else {
			pdf = 0;
		}
		out.d = d;
	} else if (em.type == MTSGPU_EMITTER_POINT || em.type == MTSGPU_EMITTER_SPOT) {
		// PointLuminaire::sample (point.cpp:58-66) / SpotLuminaire::sample (spot.cpp:117-125)
		const V3 pos = v3(em.p[0], em.p[1], em.p[2]);
		const V3 lumToP = p - pos;
		const float invDist = 1.0f / length(lumToP);
		out.p = pos; out.n = v3(0, 0, 0);
		out.d = lumToP * invDist;
		pdf = 1.0f; out.mis = false;
		if (em.type == MTSGPU_EMITTER_SPOT) {
			// falloffCurve (spot.cpp:85-109), constant texture
			const float cosTheta = em.p[3] * out.d.x + em.p[4] * out.d.y + em.p[5] * out.d.z;
			if (cosTheta <= em.p[9]) value = v3(0, 0, 0);
			else if (!(cosTheta >= em.p[8])) value = value * ((em.p[6] - mts_acos(cosTheta)) * em.p[7]);
		}
		value = value * (invDist * invDist);
	} else if (em.type == MTSGPU_EMITTER_DIRECTIONAL) {
		// DirectionalLuminaire::sample (directional.cpp:83-90)
		const V3 dir = v3(em.p[0], em.p[1], em.p[2]);
		out.p = p - dir * (2 * em.p[3]); out.n = v3(0, 0, 0);
		out.d = dir;
		pdf = 1.0f; out.mis = false;
	} else if (em.type == MTSGPU_EMITTER_COLLIMATED) {
		const BeamSample bs = collimatedSample(sc.emitterXf + (size_t) index * 32, em.p[0], p);
		out.mis = false; out.n = v3(0, 0, 0);
		if (bs.ok) { out.p = bs.p; out.d = bs.d; pdf = 1.0f; }
	} else if (em.type == MTSGPU_EMITTER_ENVMAP) {
		// EnvMapLuminaire::sample (envmap.cpp:170-182)
		const EnvSample es = envSampleDirection(mipRef(sc), sc.textures[__float_as_uint(em.p[5])], sc.envData, em.p[4], sx, sy);
		const V3 d = es.d; const float dirPdf = es.pdf; value = es.value;
		const V3 center = v3(em.p[0], em.p[1], em.p[2]);
		float nearHit, farHit;
		if (length(p - center) <= em.p[3] && bsphereRayIntersect(center, em.p[3], p, -d, nearHit, farHit)) {
			out.p = p - d * nearHit;
			out.n = normalize(center - out.p);
			out.d = d;
			pdf = dirPdf;
		}
	} else {
		// ConstantLuminaire::sample (constant.cpp:72-86)
		const V3 d = squareToSphere(sx, sy);
		const V3 center = v3(em.p[0], em.p[1], em.p[2]);
		float nearHit, farHit;
		if (length(p - center) <= em.p[3] && bsphereRayIntersect(center, em.p[3], p, d, nearHit, farHit)) {
			out.p = p + d * nearHit;
			pdf = 1.0f / (4 * MTS_PI);
			out.n = normalize(center - out.p);
			out.d = -d;
		}
	}
	if (pdf != 0) {
		pdf *= lumPdf;
		out.value = value * (1.0f / pdf);
		out.pdf = pdf;
	}
}

// Scene::pdfLuminaire (scene.cpp:381-394) with AreaLuminaire::pdf -> Shape::pdfSolidAngle (shape.cpp:77-83)
__device__ inline float pdfLuminaire(const DeviceScene &sc, V3 from, V3 lumP, V3 lumN, int emitter, bool importance) {
	const EmitterDev em = sc.emitters[emitter];
	const float luminance = importance ? em.weight : 1.0f;
	const float fraction = luminance / sc.lumOriginalSum;
	const ShapeInfo si = sc.shapes[em.shape];
	if (si.isSphere) {
		const float4 cr = __ldg(sc.positions + si.vertOffset);
		return spherePdfSolidAngle(v3(cr), cr.w, si.invSurfaceArea, from, lumP, lumN) * fraction;
	}
	const V3 lumToPoint = from - lumP;
	const float distSquared = dot(lumToPoint, lumToPoint);
	const float invDP = fmaxf(0.0f, sqrtf(distSquared) / dot(lumToPoint, lumN));
	return (si.invSurfaceArea * distSquared * invDP) * fraction;
}

// Scene::pdfLuminaire for the background luminaire: ConstantLuminaire::pdf (constant.cpp:88-90)
// `dir`: the direction of the ray that left the scene (-lRec.d)
__device__ inline float pdfBackground(const DeviceScene &sc, bool importance, V3 dir) {
	const EmitterDev em = sc.emitters[sc.background];
	const float luminance = importance ? em.weight : 1.0f;
	const float fraction = luminance / sc.lumOriginalSum;
	if (em.type == MTSGPU_EMITTER_ENVMAP) return envPdf(sc.envData, dir) * fraction;
	return (1.0f / (4 * MTS_PI)) * fraction;
}

// Scene::LeBackground (scene.h:403-405) -> ConstantLuminaire::Le (constant.cpp:68-70) or the environment map
// EnvMapLuminaire::Le(ray) = Le(normalize(ray.d)) (envmap.cpp:166-168)
__device__ inline V3 backgroundLe(const DeviceScene &sc, V3 rayD) {
	if (sc.background < 0) return v3(0, 0, 0);
	const EmitterDev em = sc.emitters[sc.background];
	if (em.type == MTSGPU_EMITTER_ENVMAP) return envLe(mipRef(sc), sc.textures[__float_as_uint(em.p[5])], sc.envData, em.p[4], normalize(rayD));
	return v3(em.intensity[0], em.intensity[1], em.intensity[2]);
}

// AreaLuminaire::Le (area.cpp:62-66)
__device__ inline V3 emitterLe(const DeviceScene &sc, int emitter, V3 n, V3 d) {
	if (dot(d, n) <= 0) return v3(0, 0, 0);
	const EmitterDev em = sc.emitters[emitter];
	return v3(em.intensity[0], em.intensity[1], em.intensity[2]);
}

// BSDF sample / eval / pdf device functions for the five models on the hot path, each restating
// the reference's arithmetic in its operation order (this TU is compiled with --fmad=false so that
// intermediate values match the reference's SSE build):
//   lambertian  src/bsdfs/lambertian.cpp:95-126      mirror      src/bsdfs/mirror.cpp:61-94
//   microfacet  src/bsdfs/microfacet.cpp:95-269      roughmetal  src/bsdfs/roughmetal.cpp:75-167
//   dielectric  src/bsdfs/dielectric.cpp:101-290     base 3-arg fallback src/librender/bsdf.cpp:37-48
//   phong       src/bsdfs/phong.cpp:104-234          roughglass  src/bsdfs/roughglass.cpp:209-598
//   ward        src/bsdfs/ward.cpp:145-285
// All directions are in the local shading frame (wi.z = cos theta). Queries use the full type
// mask, component = -1 and quantity = ERadiance, which is what path.cpp / direct.cpp pass.
#pragma once
#include "types.cuh"

#define MTS_INV_TWOPI 0.15915494309189533577f
#define BSDF_DIFFUSE_REFLECTION  0x0001u
#define BSDF_DIFFUSE_TRANSMISSION 0x0002u
#define BSDF_DELTA_REFLECTION    0x0004u
#define BSDF_DELTA_TRANSMISSION  0x0008u
#define BSDF_GLOSSY_REFLECTION   0x0010u
#define BSDF_GLOSSY_TRANSMISSION 0x0020u
#define BSDF_DELTA               (BSDF_DELTA_REFLECTION | BSDF_DELTA_TRANSMISSION)
#define BSDF_TRANSMISSION        (0x0002u | BSDF_DELTA_TRANSMISSION | BSDF_GLOSSY_TRANSMISSION)

// ---- src/libcore/util.cpp helpers ---------------------------------------------------------------
__device__ inline V3 sphericalDirection(float theta, float phi) {           // util.cpp:543-550
	float sinTheta, cosTheta, sinPhi, cosPhi;
	mts_sincos(theta, sinTheta, cosTheta); mts_sincos(phi, sinPhi, cosPhi);
	return v3(sinTheta * cosPhi, sinTheta * sinPhi, cosTheta);
}

__device__ inline V3 squareToHemispherePSA(float sx, float sy) {            // util.cpp:572-588
	const float r = sqrtf(sx);
	const float phi = 2.0f * MTS_PI * sy;
	float sinPhi, cosPhi; mts_sincos(phi, sinPhi, cosPhi);
	const float dirX = r * cosPhi;
	const float dirY = r * sinPhi;
	const float z = sqrtf(1 - fminf(1.0f, dirX * dirX + dirY * dirY));
	if (z == 0)
		return normalize(v3(dirX, dirY, MTS_EPSILON));
	return v3(dirX, dirY, z);
}

__device__ inline float fresnelDielectric(float cosTheta1, float cosTheta2, float etaI, float etaT) {   // util.cpp:680-689
	const float Rs = (etaI * cosTheta1 - etaT * cosTheta2) / (etaI * cosTheta1 + etaT * cosTheta2);
	const float Rp = (etaT * cosTheta1 - etaI * cosTheta2) / (etaT * cosTheta1 + etaI * cosTheta2);
	return (Rs * Rs + Rp * Rp) / 2.0f;
}

__device__ inline V3 fresnelConductor(float cosTheta, V3 eta, V3 k) {       // util.cpp:691-704
	const float c2 = cosTheta * cosTheta;
	const float two = 2.0f * cosTheta;
	V3 out;
	const float e[3] = { eta.x, eta.y, eta.z }, kk[3] = { k.x, k.y, k.z };
	float res[3];
	for (int i = 0; i < 3; ++i) {
		const float tmpF = e[i] * e[i] + kk[i] * kk[i];
		const float tmp = tmpF * c2;
		const float rParl2 = (tmp - (e[i] * two) + 1.0f) / (tmp + (e[i] * two) + 1.0f);
		const float rPerp2 = (tmpF - (e[i] * two) + c2) / (tmpF + (e[i] * two) + c2);
		res[i] = (rParl2 + rPerp2) / 2.0f;
	}
	out = v3(res[0], res[1], res[2]);
	return out;
}

__device__ inline float fresnel(float cosThetaI, float etaExt, float etaInt) {   // util.cpp:706-725
	float etaI = etaExt, etaT = etaInt;
	if (cosThetaI < 0.0f) { const float t = etaI; etaI = etaT; etaT = t; }
	const float sinThetaT = etaI / etaT * sqrtf(fmaxf(0.0f, 1.0f - cosThetaI * cosThetaI));
	if (sinThetaT > 1.0f)
		return 1.0f;
	const float cosThetaT = sqrtf(1.0f - sinThetaT * sinThetaT);
	return fresnelDielectric(fabsf(cosThetaI), cosThetaT, etaI, etaT);
}

__device__ inline float frameTanTheta(V3 v) {                                // core/frame.h:102-107
	const float temp = 1 - v.z * v.z;
	if (temp <= 0.0f) return 0.0f;
	return sqrtf(temp) / v.z;
}

// ---- Beckmann microfacet terms shared by microfacet.cpp:95-134 and roughmetal.cpp:75-114 -----------
__device__ inline float beckmannD(V3 m, float alphaB) {
	const float ex = frameTanTheta(m) / alphaB;
	return mts_exp(-(ex * ex)) / (MTS_PI * alphaB * alphaB * mts_pow4(m.z));
}

__device__ inline V3 sampleBeckmannD(float sx, float sy, float alphaB) {
	const float thetaM = mts_atan(sqrtf(-alphaB * alphaB * mts_log(1.0f - sx)));
	const float phiM = (2.0f * MTS_PI) * sy;
	return sphericalDirection(thetaM, phiM);
}

__device__ inline float smithBeckmannG1(V3 v, V3 m, float alphaB) {
	if (dot(v, m) * v.z <= 0)
		return 0.0f;
	const float tanTheta = frameTanTheta(v);
	if (tanTheta == 0.0f)
		return 1.0f;
	const float a = 1.0f / (alphaB * tanTheta);
	const float aSqr = a * a;
	if (a >= 1.6f)
		return 1.0f;
	return (3.535f * a + 2.181f * aSqr) / (1.0f + 2.276f * a + 2.577f * aSqr);
}

__device__ inline V3 reflectAbout(V3 wi, V3 n) {                             // n*(2*dot(n,wi)) - wi
	return n * (2.0f * dot(n, wi)) - wi;
}

__device__ inline V3 mp3(const MaterialDev &m, int o) { return v3(m.p[o], m.p[o + 1], m.p[o + 2]); }


// ---- roughglass (src/bsdfs/roughglass.cpp): Walter et al. microfacet refraction, three distributions ---------
// p[0..2] specularReflectance, p[3..5] specularTransmittance, p[6] value of the alpha texture (for the Phong
// distribution: the exponent the constructor derives, roughglass.cpp:132-138), p[7] intIOR, p[8] extIOR,
// p[9] distribution (0 beckmann, 1 phong, 2 ggx). bRec.sampler is NULL in path.cpp / direct.cpp
// (records.inl:24-37), so the branch that recycles the 2D sample is the one that runs.
__device__ inline float rgSignum(float v) { return (v < 0) ? -1.0f : 1.0f; }                 // roughglass.cpp:198-200

__device__ inline float rgAlpha(const MaterialDev &m) {          // ConstantFloatTexture -> Spectrum::average (spectrum.h:299-304)
	float result = 0.0f;
	result += m.p[6]; result += m.p[6]; result += m.p[6];
	return result * (1.0f / 3);
}

__device__ inline float rgEvalD(V3 m, float alpha, int distr) {                                // roughglass.cpp:209-252
	if (m.z <= 0) return 0.0f;
	float result;
	if (distr == 0) {
		const float ex = frameTanTheta(m) / alpha;
		result = mts_exp(-(ex * ex)) / (MTS_PI * alpha * alpha * mts_pow4(m.z));
	} else if (distr == 1) {
		result = (alpha + 2) * MTS_INV_TWOPI * mts_pow(m.z, alpha);
	} else {
		const float tanTheta = frameTanTheta(m), cosTheta = m.z;
		const float root = alpha / (cosTheta * cosTheta * (alpha * alpha + tanTheta * tanTheta));
		result = MTS_INV_PI * (root * root);
	}
	if ((double) result < 1e-40) result = 0;
	return result;
}

__device__ inline V3 rgSampleD(float sx, float sy, float alpha, int distr) {                    // roughglass.cpp:261-288
	const float phiM = (2.0f * MTS_PI) * sy;
	float thetaM;
	if (distr == 0) thetaM = mts_atan(sqrtf(-alpha * alpha * mts_log(1.0f - sx)));
	else if (distr == 1) thetaM = mts_acos(mts_pow(sx, (float) 1 / (alpha + 2)));
	else thetaM = mts_atan(alpha * sqrtf(sx) / sqrtf(1.0f - sx));
	return sphericalDirection(thetaM, phiM);
}

__device__ inline float rgSmithG1(V3 v, V3 m, float alpha, int distr) {                         // roughglass.cpp:298-340
	const float tanTheta = fabsf(frameTanTheta(v));
	if (tanTheta == 0.0f) return 1.0f;
	if (dot(v, m) * v.z <= 0) return 0.0f;
	if (distr == 2) {
		const float root = alpha * tanTheta;
		return 2.0f / (1.0f + sqrtf(1.0f + root * root));
	}
	if (distr == 1) alpha = sqrtf(0.5f * alpha + 1) / tanTheta;
	const float a = 1.0f / (alpha * tanTheta);
	const float aSqr = a * a;
	if (a >= 1.6f) return 1.0f;
	return (3.535f * a + 2.181f * aSqr) / (1.0f + 2.276f * a + 2.577f * aSqr);
}

// half-vector of a reflection / transmission pair (roughglass.cpp:356-376)
__device__ inline V3 rgHalf(V3 wi, V3 wo, bool reflect, float etaI, float etaT, float extIOR, float intIOR) {
	if (reflect) return normalize(wo + wi) * rgSignum(wo.z);
	return normalize(wi * etaI + wo * etaT) * (extIOR > intIOR ? 1.0f : -1.0f);
}

__device__ inline V3 roughglassF(const MaterialDev &m, V3 wi, V3 wo) {                         // roughglass.cpp:345-412
	const float intIOR = m.p[7], extIOR = m.p[8]; const int distr = (int) m.p[9];
	const bool reflect = wi.z * wo.z > 0;
	float etaI = extIOR, etaT = intIOR;
	if (wi.z < 0) { const float t = etaI; etaI = etaT; etaT = t; }
	const V3 H = rgHalf(wi, wo, reflect, etaI, etaT, extIOR, intIOR);
	const float alpha = rgAlpha(m);
	const float D = rgEvalD(H, alpha, distr);
	if (D == 0) return v3(0, 0, 0);
	const float F = fresnel(dot(wi, H), extIOR, intIOR);
	const float G = rgSmithG1(wi, H, alpha, distr) * rgSmithG1(wo, H, alpha, distr);
	if (reflect) {
		const float value = F * D * G / (4.0f * wi.z * wo.z);
		return mp3(m, 0) * value;
	}
	const float sqrtDenom = etaI * dot(wi, H) + etaT * dot(wo, H);
	float value = ((1 - F) * D * G * etaT * etaT * dot(wi, H) * dot(wo, H)) / (wi.z * wo.z * sqrtDenom * sqrtDenom);
	value *= (etaI * etaI) / (etaT * etaT);            // quantity == ERadiance
	return mp3(m, 3) * fabsf(value);
}

__device__ inline float rgClampedFresnel(const MaterialDev &m, V3 wi) {
	return fminf(0.9f, fmaxf(0.1f, fresnel(wi.z, m.p[8], m.p[7])));
}

__device__ inline float roughglassPdf(const MaterialDev &m, V3 wi, V3 wo) {                    // roughglass.cpp:414-488
	const float intIOR = m.p[7], extIOR = m.p[8]; const int distr = (int) m.p[9];
	const bool reflect = wi.z * wo.z > 0;
	float etaI = extIOR, etaT = intIOR;
	if (wi.z < 0) { const float t = etaI; etaI = etaT; etaT = t; }
	const V3 H = rgHalf(wi, wo, reflect, etaI, etaT, extIOR, intIOR);
	float dwh_dwo;
	if (reflect) {
		dwh_dwo = 1.0f / (4.0f * dot(wo, H));
	} else {
		const float sqrtDenom = etaI * dot(wi, H) + etaT * dot(wo, H);
		dwh_dwo = (etaT * etaT * dot(wo, H)) / (sqrtDenom * sqrtDenom);
	}
	float alpha = rgAlpha(m);
	alpha = alpha * (1.2f - 0.2f * sqrtf(fabsf(wi.z)));
	float prob = rgEvalD(H, alpha, distr);
	const float F = rgClampedFresnel(m, wi);
	prob *= reflect ? F : (1 - F);
	return fabsf(prob * H.z * dwh_dwo);
}

// 2-argument sample (roughglass.cpp:490-598): returns the sampling weight; zero => no sample
__device__ inline V3 roughglassSample2(const MaterialDev &m, V3 wi, float sx, float sy, V3 &wo, uint32_t &type) {
	const V3 zero = v3(0, 0, 0);
	const float intIOR = m.p[7], extIOR = m.p[8]; const int distr = (int) m.p[9];
	bool choseReflection = true;
	float sampleF = rgClampedFresnel(m, wi);
	if (sx < sampleF) {
		sx /= sampleF;
	} else {
		sx = (sx - sampleF) / (1 - sampleF);
		choseReflection = false;
	}
	const float alpha = rgAlpha(m);
	const float sampleAlpha = alpha * (1.2f - 0.2f * sqrtf(fabsf(wi.z)));
	const V3 mN = rgSampleD(sx, sy, sampleAlpha, distr);
	V3 result;
	if (choseReflection) {
		wo = mN * (2 * dot(wi, mN)) - wi;
		type = BSDF_GLOSSY_REFLECTION;
		if (wi.z * wo.z <= 0) return zero;
		result = mp3(m, 0);
	} else {
		float etaI = extIOR, etaT = intIOR;
		if (wi.z < 0) { const float t = etaI; etaI = etaT; etaT = t; }
		const float eta = etaI / etaT, c = dot(wi, mN);                     // refract(), roughglass.cpp:181-196
		const float cosThetaTSqr = 1 + eta * eta * (c * c - 1);
		if (cosThetaTSqr < 0) return zero;
		wo = mN * (eta * c - rgSignum(wi.z) * sqrtf(cosThetaTSqr)) - wi * eta;
		type = BSDF_GLOSSY_TRANSMISSION;
		if (wi.z * wo.z >= 0) return zero;
		result = mp3(m, 3) * ((etaI * etaI) / (etaT * etaT));
	}
	float numerator = rgEvalD(mN, alpha, distr) * rgSmithG1(wi, mN, alpha, distr) * rgSmithG1(wo, mN, alpha, distr) * dot(wi, mN);
	float denominator = rgEvalD(mN, sampleAlpha, distr) * mN.z * wi.z * wo.z;
	float F = fresnel(dot(wi, mN), extIOR, intIOR);
	if (!choseReflection) { sampleF = 1 - sampleF; F = 1 - F; }
	numerator *= F;
	denominator *= sampleF;
	return result * fabsf(numerator / denominator);
}

// ---- ward (src/bsdfs/ward.cpp): p[6] kd, p[7] ks, p[8] alphaX, p[9] alphaY, p[10] / p[11] sampling weights, p[12] model ----
// std::pow(Float, int) promotes to double (C++11), and the reference keeps that double through the rest of the
// expression before rounding to Float; the two factors below do the same.
__device__ inline float wardSpecRef(const MaterialDev &m, V3 wi, V3 wo) {                    // ward.cpp:157-186
	const float ax = m.p[8], ay = m.p[9];
	const V3 H = wi + wo;
	float factor1 = 0.0f;
	const int model = (int) m.p[12];
	if (model == 0) factor1 = 1.0f / (4.0f * MTS_PI * ax * ay * sqrtf(wi.z * wo.z));
	else if (model == 1) factor1 = 1.0f / (4.0f * MTS_PI * ax * ay * wi.z * wo.z);
	else factor1 = (float) ((double) dot(H, H) / ((double) (MTS_PI * ax * ay) * mts_pow4d(H.z)));
	const float factor2 = H.x / ax, factor3 = H.y / ay;
	const float exponent = -(factor2 * factor2 + factor3 * factor3) / (H.z * H.z);
	return factor1 * mts_exp(exponent) * m.p[7];
}
__device__ inline float wardPdfSpec(const MaterialDev &m, V3 wi, V3 wo) {                    // ward.cpp:196-205
	const float ax = m.p[8], ay = m.p[9];
	const V3 H = normalize(wi + wo);
	const float factor1 = (float) (1.0 / ((double) (4.0f * MTS_PI * ax * ay * dot(H, wi)) * ((double) H.z * (double) H.z * (double) H.z)));
	const float factor2 = H.x / ax, factor3 = H.y / ay;
	const float exponent = -(factor2 * factor2 + factor3 * factor3) / (H.z * H.z);
	return factor1 * mts_exp(exponent);
}
__device__ inline V3 wardF(const MaterialDev &m, V3 wi, V3 wo) {                             // ward.cpp:145-194
	V3 result = v3(0, 0, 0);
	if (wi.z <= 0 || wo.z <= 0) return result;
	const float specRef = wardSpecRef(m, wi, wo);
	if (specRef > 1e-10f) result = result + mp3(m, 3) * specRef;
	result = result + mp3(m, 0) * (MTS_INV_PI * m.p[6]);
	return result;
}
__device__ inline float wardPdf(const MaterialDev &m, V3 wi, V3 wo) {                        // ward.cpp:207-227
	if (wi.z <= 0 || wo.z <= 0) return 0.0f;
	return m.p[10] * wardPdfSpec(m, wi, wo) + m.p[11] * (wo.z * MTS_INV_PI);
}
// half vector of the specular lobe (ward.cpp:229-243)
__device__ inline V3 wardSampleH(const MaterialDev &m, float sx, float sy) {
	const float ax = m.p[8], ay = m.p[9];
	float phiH = mts_atan(ay / ax * mts_tan(2.0f * MTS_PI * sy));
	if (sy > 0.5f) phiH += MTS_PI;
	const float cosPhiH = mts_cos(phiH);
	const float sinPhiH = sqrtf(fmaxf(0.0f, 1.0f - cosPhiH * cosPhiH));
	const float thetaH = mts_atan(sqrtf(fmaxf(0.0f, -mts_log(sx) / ((cosPhiH * cosPhiH) / (ax * ax) + (sinPhiH * sinPhiH) / (ay * ay)))));
	return sphericalDirection(thetaH, phiH);
}

// ---- f ---------------------------------------------------------------------------------------------
__device__ inline V3 bsdfF_base(const MaterialDev &m, V3 wi, V3 wo) {
	const V3 zero = v3(0, 0, 0);
	switch (m.type & MTSGPU_BSDF_TYPE_MASK) {
		case MTSGPU_BSDF_LAMBERTIAN:
			if (wi.z <= 0 || wo.z <= 0) return zero;
			return mp3(m, 0) * MTS_INV_PI;
		case MTSGPU_BSDF_DIFFTRANS:                         // difftrans.cpp:91-96
			if (wi.z * wo.z >= 0) return zero;
			return mp3(m, 0) * MTS_INV_PI;
		case MTSGPU_BSDF_MICROFACET: {
			if (wi.z <= 0 || wo.z <= 0) return zero;
			const float kd = m.p[6], ks = m.p[7], alphaB = m.p[8], intIOR = m.p[9], extIOR = m.p[10];
			const V3 Hr = normalize(wi + wo);
			const float F = fresnel(dot(wi, Hr), extIOR, intIOR);
			const float D = beckmannD(Hr, alphaB);
			const float G = smithBeckmannG1(wi, Hr, alphaB) * smithBeckmannG1(wo, Hr, alphaB);
			const float specRef = D * G / (4.0f * wi.z * wo.z);
			V3 result = (mp3(m, 3) * specRef) * (F * ks);
			result = result + mp3(m, 0) * (MTS_INV_PI * (1 - F) * kd);
			return result;
		}
		case MTSGPU_BSDF_ROUGHMETAL: {
			if (wi.z <= 0 || wo.z <= 0) return zero;
			const float alphaB = m.p[3];
			const V3 Hr = normalize(wi + wo);
			const V3 F = fresnelConductor(dot(wi, Hr), mp3(m, 4), mp3(m, 7));
			const float D = beckmannD(Hr, alphaB);
			const float G = smithBeckmannG1(wi, Hr, alphaB) * smithBeckmannG1(wo, Hr, alphaB);
			const V3 specRef = F * (D * G / (4.0f * wi.z * wo.z));
			return mp3(m, 0) * specRef;
		}
		case MTSGPU_BSDF_PHONG: {                  // phong.cpp:104-131
			if (wi.z <= 0 || wo.z <= 0) return zero;
			const float kd = m.p[6], ks = m.p[7], exponent = m.p[8];
			const float alpha = dot(v3(-wi.x, -wi.y, wi.z), wo);
			float specRef;
			if (alpha <= 0.0f) specRef = 0.0f;
			else specRef = (exponent + 2) * MTS_INV_TWOPI * mts_pow(alpha, exponent) * ks;
			V3 result = mp3(m, 3) * specRef;
			result = result + mp3(m, 0) * (MTS_INV_PI * kd);
			return result;
		}
		case MTSGPU_BSDF_ROUGHGLASS:
			return roughglassF(m, wi, wo);
		case MTSGPU_BSDF_WARD:
			return wardF(m, wi, wo);
		default:   // mirror, dielectric: delta lobes only
			return zero;
	}
}

// ---- pdf -------------------------------------------------------------------------------------------
__device__ inline float pdfSpecBeckmann(V3 wi, V3 wo, float alphaB) {
	const V3 Hr = normalize(wi + wo);
	return beckmannD(Hr, alphaB) * Hr.z / (4.0f * fabsf(dot(wo, Hr)));
}

__device__ inline float bsdfPdf_base(const MaterialDev &m, V3 wi, V3 wo) {
	switch (m.type & MTSGPU_BSDF_TYPE_MASK) {
		case MTSGPU_BSDF_LAMBERTIAN:
			if (wi.z <= 0 || wo.z <= 0) return 0.0f;
			return wo.z * MTS_INV_PI;
		case MTSGPU_BSDF_DIFFTRANS:                         // difftrans.cpp:98-102
			if (wi.z * wo.z >= 0) return 0.0f;
			return fabsf(wo.z) * MTS_INV_PI;
		case MTSGPU_BSDF_MICROFACET: {
			if (wi.z <= 0 || wo.z <= 0) return 0.0f;
			const float kd = m.p[6], ks = m.p[7], alphaB = m.p[8], intIOR = m.p[9], extIOR = m.p[10];
			float fr = fresnel(wi.z, extIOR, intIOR);
			fr = fminf(fmaxf(fr, 0.05f), 0.95f);
			const float diffuseSamplingWeight = (1 - fr) * kd;
			const float specularSamplingWeight = fr * ks;
			const float normalization = 1 / (diffuseSamplingWeight + specularSamplingWeight);
			return (specularSamplingWeight * pdfSpecBeckmann(wi, wo, alphaB)
				+ diffuseSamplingWeight * (wo.z * MTS_INV_PI)) * normalization;
		}
		case MTSGPU_BSDF_ROUGHMETAL: {
			if (wi.z <= 0 || wo.z <= 0) return 0.0f;
			const float alphaB = m.p[3];
			const V3 Hr = normalize(wi + wo);
			const float dwhr_dwo = 1.0f / (4.0f * fabsf(dot(wo, Hr)));
			return beckmannD(Hr, alphaB) * Hr.z * dwhr_dwo;
		}
		case MTSGPU_BSDF_PHONG: {                  // phong.cpp:133-162
			if (wo.z <= 0 || wi.z <= 0) return 0.0f;
			const float exponent = m.p[8];
			const float alpha = dot(v3(-wi.x, -wi.y, wi.z), wo);
			float specPdf = mts_pow(alpha, exponent) * (exponent + 1.0f) / (2.0f * MTS_PI);
			if (alpha <= 0) specPdf = 0;
			return m.p[9] * specPdf + m.p[10] * (wo.z * MTS_INV_PI);
		}
		case MTSGPU_BSDF_ROUGHGLASS:
			return roughglassPdf(m, wi, wo);
		case MTSGPU_BSDF_WARD:
			return wardPdf(m, wi, wo);
		default:
			return 0.0f;
	}
}

// ---- f and pdf together ----------------------------------------------------------------------------
// The reference evaluates f() and pdf() separately (and again after sampling, bsdf.cpp:37-48); both depend on
// the same half-vector and Beckmann term. Computing those once gives bit-identical values (same expressions on
// the same inputs) for a third of the transcendental work.
__device__ inline void bsdfFPdf_base(const MaterialDev &m, V3 wi, V3 wo, V3 &f, float &pdf) {
	f = v3(0, 0, 0); pdf = 0.0f;
	switch (m.type & MTSGPU_BSDF_TYPE_MASK) {
		case MTSGPU_BSDF_LAMBERTIAN:
			if (wi.z <= 0 || wo.z <= 0) return;
			f = mp3(m, 0) * MTS_INV_PI; pdf = wo.z * MTS_INV_PI;
			return;
		case MTSGPU_BSDF_DIFFTRANS:
			if (wi.z * wo.z >= 0) return;
			f = mp3(m, 0) * MTS_INV_PI; pdf = fabsf(wo.z) * MTS_INV_PI;
			return;
		case MTSGPU_BSDF_MICROFACET: {
			if (wi.z <= 0 || wo.z <= 0) return;
			const float kd = m.p[6], ks = m.p[7], alphaB = m.p[8], intIOR = m.p[9], extIOR = m.p[10];
			const V3 Hr = normalize(wi + wo);
			const float F = fresnel(dot(wi, Hr), extIOR, intIOR);
			const float D = beckmannD(Hr, alphaB);
			const float G = smithBeckmannG1(wi, Hr, alphaB) * smithBeckmannG1(wo, Hr, alphaB);
			const float specRef = D * G / (4.0f * wi.z * wo.z);
			f = (mp3(m, 3) * specRef) * (F * ks);
			f = f + mp3(m, 0) * (MTS_INV_PI * (1 - F) * kd);
			float fr = fresnel(wi.z, extIOR, intIOR);
			fr = fminf(fmaxf(fr, 0.05f), 0.95f);
			const float diffuseSamplingWeight = (1 - fr) * kd;
			const float specularSamplingWeight = fr * ks;
			const float normalization = 1 / (diffuseSamplingWeight + specularSamplingWeight);
			const float pdfSpec = D * Hr.z / (4.0f * fabsf(dot(wo, Hr)));
			pdf = (specularSamplingWeight * pdfSpec + diffuseSamplingWeight * (wo.z * MTS_INV_PI)) * normalization;
			return;
		}
		case MTSGPU_BSDF_ROUGHMETAL: {
			if (wi.z <= 0 || wo.z <= 0) return;
			const float alphaB = m.p[3];
			const V3 Hr = normalize(wi + wo);
			const V3 F = fresnelConductor(dot(wi, Hr), mp3(m, 4), mp3(m, 7));
			const float D = beckmannD(Hr, alphaB);
			const float G = smithBeckmannG1(wi, Hr, alphaB) * smithBeckmannG1(wo, Hr, alphaB);
			f = mp3(m, 0) * (F * (D * G / (4.0f * wi.z * wo.z)));
			const float dwhr_dwo = 1.0f / (4.0f * fabsf(dot(wo, Hr)));
			pdf = D * Hr.z * dwhr_dwo;
			return;
		}
		case MTSGPU_BSDF_PHONG:
		case MTSGPU_BSDF_ROUGHGLASS:
		case MTSGPU_BSDF_WARD:
			f = bsdfF_base(m, wi, wo); pdf = bsdfPdf_base(m, wi, wo);
			return;
		default:
			return;
	}
}

// Sampling of the two rough models (microfacet.cpp:201-269, roughmetal.cpp:147-167): direction, then f and pdf at
// it. Returns false exactly when the reference's 2-argument sample() returns a zero spectrum.
__device__ inline bool bsdfSampleRough(const MaterialDev &m, V3 wi, float sx, float sy, V3 &wo, uint32_t &type, V3 &f, float &pdf, V3 &ratio) {
	if (wi.z <= 0) return false;
	if ((m.type & MTSGPU_BSDF_TYPE_MASK) == MTSGPU_BSDF_MICROFACET) {
		const float kd = m.p[6], ks = m.p[7], alphaB = m.p[8], intIOR = m.p[9], extIOR = m.p[10];
		float fr = fresnel(wi.z, extIOR, intIOR);
		fr = fminf(fmaxf(fr, 0.05f), 0.95f);
		float diffuseSamplingWeight = (1 - fr) * kd;
		float specularSamplingWeight = fr * ks;
		const float normalization = 1 / (diffuseSamplingWeight + specularSamplingWeight);
		specularSamplingWeight *= normalization;
		diffuseSamplingWeight *= normalization;
		if (sx < specularSamplingWeight) {
			sx /= specularSamplingWeight;
			wo = reflectAbout(wi, sampleBeckmannD(sx, sy, alphaB));
			type = BSDF_GLOSSY_REFLECTION;
			if (wo.z <= 0) return false;
			bsdfFPdf_base(m, wi, wo, f, pdf);
			if (pdf == 0) return false;
		} else {
			sx = (sx - specularSamplingWeight) / diffuseSamplingWeight;
			wo = squareToHemispherePSA(sx, sy);
			type = BSDF_DIFFUSE_REFLECTION;
			bsdfFPdf_base(m, wi, wo, f, pdf);
		}
	} else if ((m.type & MTSGPU_BSDF_TYPE_MASK) == MTSGPU_BSDF_PHONG) {      // phong.cpp:164-234
		const float exponent = m.p[8], ssw = m.p[9], dsw = m.p[10];
		if (sx <= ssw) {
			sx /= ssw;
			const V3 R = v3(-wi.x, -wi.y, wi.z);
			const float sinAlpha = sqrtf(1 - mts_pow(sy, 2 / (exponent + 1)));
			const float cosAlpha = mts_pow(sy, 1 / (exponent + 1));
			const float phi = (2.0f * MTS_PI) * sx;
			const V3 localDir = v3(sinAlpha * mts_cos(phi), sinAlpha * mts_sin(phi), cosAlpha);
			wo = toWorld(makeFrame(R), localDir);
			type = BSDF_GLOSSY_REFLECTION;
			if (wo.z <= 0) return false;
			bsdfFPdf_base(m, wi, wo, f, pdf);
			if (pdf == 0) return false;
		} else {
			sx = (sx - ssw) / dsw;
			wo = squareToHemispherePSA(sx, sy);
			type = BSDF_DIFFUSE_REFLECTION;
			bsdfFPdf_base(m, wi, wo, f, pdf);
		}
	} else if ((m.type & MTSGPU_BSDF_TYPE_MASK) == MTSGPU_BSDF_WARD) {       // ward.cpp:229-285
		const float ssw = m.p[10], dsw = m.p[11];
		if (sx <= ssw) {
			sx = sx / ssw;
			const V3 H = wardSampleH(m, sx, sy);
			wo = H * (2.0f * dot(wi, H)) - wi;
			type = BSDF_GLOSSY_REFLECTION;
			if (wo.z <= 0.0f) return false;
		} else {
			sx = (sx - ssw) / dsw;
			wo = squareToHemispherePSA(sx, sy);
			type = BSDF_DIFFUSE_REFLECTION;
		}
		bsdfFPdf_base(m, wi, wo, f, pdf);
	} else {
		wo = reflectAbout(wi, sampleBeckmannD(sx, sy, m.p[3]));
		type = BSDF_GLOSSY_REFLECTION;
		if (wo.z <= 0) return false;
		bsdfFPdf_base(m, wi, wo, f, pdf);
	}
	ratio = f * (1.0f / pdf);
	return !isZero(ratio);
}

// ---- dielectric core (dielectric.cpp:116-146): Fresnel reflectance + transmitted cosine -------------
__device__ inline void dielectricSetup(const MaterialDev &m, V3 wi, float &Fr, float &eta, float &cosThetaT) {
	const float cosThetaI = wi.z;
	float etaI = m.p[7], etaT = m.p[6];       // extIOR, intIOR
	const bool entering = cosThetaI > 0.0f;
	if (!entering) { const float t = etaI; etaI = etaT; etaT = t; }
	eta = etaI / etaT;
	const float sinThetaTSqr = eta * eta * (1.0f - wi.z * wi.z);
	cosThetaT = 0;
	if (sinThetaTSqr >= 1.0f) {
		Fr = 1.0f;
	} else {
		cosThetaT = sqrtf(1.0f - sinThetaTSqr);
		Fr = fresnelDielectric(fabsf(cosThetaI), cosThetaT, etaI, etaT);
		if (entering)
			cosThetaT = -cosThetaT;
	}
}

// ---- 2-argument sample: returns f / pdf (NOT multiplied by the cosine); zero => no sample ----------
__device__ inline V3 bsdfSample2_base(const MaterialDev &m, V3 wi, float sx, float sy, V3 &wo, uint32_t &type) {
	const V3 zero = v3(0, 0, 0);
	switch (m.type & MTSGPU_BSDF_TYPE_MASK) {
		case MTSGPU_BSDF_LAMBERTIAN:
			if (wi.z <= 0) return zero;
			wo = squareToHemispherePSA(sx, sy);
			type = BSDF_DIFFUSE_REFLECTION;
			return mp3(m, 0) * (1.0f / wo.z);              // Spectrum / Float multiplies by the reciprocal
		case MTSGPU_BSDF_DIFFTRANS:                         // difftrans.cpp:104-118
			wo = squareToHemispherePSA(sx, sy);
			if (wi.z > 0) wo.z *= -1;
			type = BSDF_DIFFUSE_TRANSMISSION;
			if (wo.z == 0) return zero;
			return mp3(m, 0) * (1.0f / fabsf(wo.z));
		case MTSGPU_BSDF_MIRROR:
			wo = v3(-wi.x, -wi.y, wi.z);
			type = BSDF_DELTA_REFLECTION;
			return mp3(m, 0);                               // Mirror::sample(2-arg) forwards to the 3-arg one
		case MTSGPU_BSDF_MICROFACET:
		case MTSGPU_BSDF_PHONG:
		case MTSGPU_BSDF_WARD:
		case MTSGPU_BSDF_ROUGHMETAL: {
			V3 f, ratio; float pdf;
			return bsdfSampleRough(m, wi, sx, sy, wo, type, f, pdf, ratio) ? ratio : zero;
		}
		case MTSGPU_BSDF_ROUGHGLASS:
			return roughglassSample2(m, wi, sx, sy, wo, type);
		case MTSGPU_BSDF_DIELECTRIC: {
			float Fr, eta, cosThetaT;
			dielectricSetup(m, wi, Fr, eta, cosThetaT);
			if (sx <= Fr) {
				type = BSDF_DELTA_REFLECTION;
				wo = v3(-wi.x, -wi.y, wi.z);
				return mp3(m, 0) * (1.0f / fabsf(wo.z));
			} else {
				type = BSDF_DELTA_TRANSMISSION;
				wo = v3(-eta * wi.x, -eta * wi.y, cosThetaT);
				return (mp3(m, 3) * (eta * eta)) * (1.0f / fabsf(wo.z));
			}
		}
		default:
			return zero;
	}
}

// ---- 3-argument sample: returns f (not divided by pdf) and the pdf; zero => no sample ---------------
__device__ inline V3 bsdfSample3_base(const MaterialDev &m, V3 wi, float sx, float sy, V3 &wo, float &pdf, uint32_t &type) {
	const V3 zero = v3(0, 0, 0);
	switch (m.type & MTSGPU_BSDF_TYPE_MASK) {
		case MTSGPU_BSDF_LAMBERTIAN:
			if (wi.z <= 0) return zero;
			wo = squareToHemispherePSA(sx, sy);
			type = BSDF_DIFFUSE_REFLECTION;
			pdf = wo.z * MTS_INV_PI;
			return mp3(m, 0) * MTS_INV_PI;
		case MTSGPU_BSDF_DIFFTRANS:                         // difftrans.cpp:120-135
			wo = squareToHemispherePSA(sx, sy);
			if (wi.z > 0) wo.z *= -1;
			type = BSDF_DIFFUSE_TRANSMISSION;
			pdf = fabsf(wo.z) * MTS_INV_PI;
			if (wo.z == 0) return zero;
			return mp3(m, 0) * MTS_INV_PI;
		case MTSGPU_BSDF_MIRROR:
			wo = v3(-wi.x, -wi.y, wi.z);
			type = BSDF_DELTA_REFLECTION;
			pdf = fabsf(wo.z);
			return mp3(m, 0);
		case MTSGPU_BSDF_DIELECTRIC: {
			float Fr, eta, cosThetaT;
			dielectricSetup(m, wi, Fr, eta, cosThetaT);
			if (sx <= Fr) {
				type = BSDF_DELTA_REFLECTION;
				wo = v3(-wi.x, -wi.y, wi.z);
				pdf = Fr * fabsf(wo.z);
				return mp3(m, 0) * Fr;
			} else {
				type = BSDF_DELTA_TRANSMISSION;
				wo = v3(-eta * wi.x, -eta * wi.y, cosThetaT);
				pdf = (1 - Fr) * fabsf(wo.z);
				return (mp3(m, 3) * (1 - Fr)) * (eta * eta);
			}
		}
		case MTSGPU_BSDF_ROUGHGLASS: {
			// RoughGlass's 3-argument sample takes the pdf by value (roughglass.cpp:619) and so never overrides
			// the virtual: the base fallback runs (bsdf.cpp:37-48)
			if (isZero(roughglassSample2(m, wi, sx, sy, wo, type))) { pdf = 0.0f; return zero; }
			pdf = roughglassPdf(m, wi, wo);
			return roughglassF(m, wi, wo);
		}
		default: {
			// BSDF::sample(bRec, pdf, sample) fallback (bsdf.cpp:37-48): 2-arg sample, then pdf() and f()
			V3 f, ratio;
			if (!bsdfSampleRough(m, wi, sx, sy, wo, type, f, pdf, ratio)) { pdf = 0.0f; return zero; }
			return f;          // pdf() and f() re-evaluated by the reference give these same values
		}
	}
}

// ---- twosided adapter (src/bsdfs/twosided.cpp:81-131): a BRDF seen from below behaves as from above --------
__device__ inline V3 bsdfF_two(const MaterialDev &m, V3 wi, V3 wo) {
	if ((m.type & MTSGPU_BSDF_TWOSIDED) && wi.z < 0) { wi.z *= -1; wo.z *= -1; }
	return bsdfF_base(m, wi, wo);
}
__device__ inline void bsdfFPdf_two(const MaterialDev &m, V3 wi, V3 wo, V3 &f, float &pdf) {
	if ((m.type & MTSGPU_BSDF_TWOSIDED) && wi.z < 0) { wi.z *= -1; wo.z *= -1; }
	bsdfFPdf_base(m, wi, wo, f, pdf);
}
__device__ inline float bsdfPdf_two(const MaterialDev &m, V3 wi, V3 wo) {
	if ((m.type & MTSGPU_BSDF_TWOSIDED) && wi.z < 0) { wi.z *= -1; wo.z *= -1; }
	return bsdfPdf_base(m, wi, wo);
}
__device__ inline V3 bsdfSample2_two(const MaterialDev &m, V3 wi, float sx, float sy, V3 &wo, uint32_t &type) {
	const bool flipped = (m.type & MTSGPU_BSDF_TWOSIDED) && wi.z < 0;
	if (flipped) wi.z *= -1;
	const V3 result = bsdfSample2_base(m, wi, sx, sy, wo, type);
	if (flipped && !isZero(result)) wo.z *= -1;
	return result;
}
__device__ inline V3 bsdfSample3_two(const MaterialDev &m, V3 wi, float sx, float sy, V3 &wo, float &pdf, uint32_t &type) {
	const bool flipped = (m.type & MTSGPU_BSDF_TWOSIDED) && wi.z < 0;
	if (flipped) wi.z *= -1;
	const V3 result = bsdfSample3_base(m, wi, sx, sy, wo, pdf, type);
	if (flipped && !isZero(result) && pdf != 0) wo.z *= -1;
	return result;
}

// ---- mask wrapper (src/bsdfs/mask.cpp:68-160): opacity-weighted nested BSDF + a delta transmission straight through ----
__device__ inline float maskOpacity(const MaterialDev &m) {     // m_opacity->getValue(its).getLuminance()
	return m.p[13] * 0.212671f + m.p[14] * 0.715160f + m.p[15] * 0.072169f;
}
__device__ inline V3 bsdfF(const MaterialDev &m, V3 wi, V3 wo) {
	const V3 f = bsdfF_two(m, wi, wo);
	return (m.type & MTSGPU_BSDF_MASKED) ? f * maskOpacity(m) : f;
}
__device__ inline void bsdfFPdf(const MaterialDev &m, V3 wi, V3 wo, V3 &f, float &pdf) {
	bsdfFPdf_two(m, wi, wo, f, pdf);
	if (m.type & MTSGPU_BSDF_MASKED) { const float o = maskOpacity(m); f = f * o; pdf = pdf * o; }
}
__device__ inline float bsdfPdf(const MaterialDev &m, V3 wi, V3 wo) {
	const float pdf = bsdfPdf_two(m, wi, wo);
	return (m.type & MTSGPU_BSDF_MASKED) ? pdf * maskOpacity(m) : pdf;
}
// (Both samplers reach the nested BSDF through ONE call site: with a call in the masked and one in the plain branch ptxas kept
// the nested sampler — 20 000 instructions — out of line, and with it the material record, wo, pdf and type in local memory.)
__device__ inline V3 bsdfSample2(const MaterialDev &m, V3 wi, float sx, float sy, V3 &wo, uint32_t &type) {
	const bool masked = (m.type & MTSGPU_BSDF_MASKED) != 0;
	if (masked) {
		const float probBSDF = maskOpacity(m);                   // mask.cpp:91-112
		if (!(sx <= probBSDF)) {
			wo = v3(-wi.x, -wi.y, -wi.z);
			type = BSDF_DELTA_TRANSMISSION;
			const float r = 1 / fabsf(wo.z);
			return v3(r, r, r);
		}
		sx = sx / probBSDF;
	}
	return bsdfSample2_two(m, wi, sx, sy, wo, type);
}
__device__ inline V3 bsdfSample3(const MaterialDev &m, V3 wi, float sx, float sy, V3 &wo, float &pdf, uint32_t &type) {
	const bool masked = (m.type & MTSGPU_BSDF_MASKED) != 0;
	float probBSDF = 1.0f;
	if (masked) {
		probBSDF = maskOpacity(m);                               // mask.cpp:125-147
		if (!(sx <= probBSDF)) {
			wo = v3(-wi.x, -wi.y, -wi.z);
			type = BSDF_DELTA_TRANSMISSION;
			pdf = (1 - probBSDF) * fabsf(wo.z);
			const float r = 1 - probBSDF;
			return v3(r, r, r);
		}
		sx = sx / probBSDF;
	}
	V3 result = bsdfSample3_two(m, wi, sx, sy, wo, pdf, type);
	if (masked) { result = result * probBSDF; pdf *= probBSDF; }
	return result;
}

/* TEST INFRASTRUCTURE — oracle harness. Not part of the product; nothing in the CUDA
 * library or the gpupath plugin calls into this file.
 *
 * Links against oracle/_ref/<variant>/libmts_ref.so (the reference's own libcore +
 * librender, compiled from /root/reference by oracle/Makefile) and drives the reference's
 * real classes the way src/mitsuba/mitsuba.cpp + src/librender/scenehandler.cpp do:
 * objects come from PluginManager::createObject, are wired with addChild/setParent/
 * configure (scenehandler.cpp:413-447), rendered by a RenderJob on N LocalWorkers
 * (mitsuba.cpp:243-244,291,364-373).
 *
 * Commands (all binary I/O is raw little-endian float32/uint32):
 *   trace   SCENE RAYS OUT closest|closest_full|any [threads] [repeat]
 *   bsdf    SCENE QUERIES OUT
 *   emitter SCENE QUERIES OUT
 *   camera  SCENE SAMPLES OUT
 *   splat   SCENE SAMPLES OUT
 *   render  SCENE OUT [threads] [integratorPlugin]
 * Each prints one JSON line with timings/counters on stdout.
 */
#include <atomic>
#include <mitsuba/render/scene.h>
#include <mitsuba/render/renderjob.h>
#include <mitsuba/render/renderqueue.h>
#include <mitsuba/render/trimesh.h>
#include <mitsuba/render/imageblock.h>
#include <mitsuba/core/plugin.h>
#include <mitsuba/core/statistics.h>
#include <mitsuba/core/fresolver.h>
#include <mitsuba/core/sched.h>
#include <mitsuba/core/appender.h>
#include <mitsuba/core/bitmap.h>
#include <mitsuba/core/fstream.h>
#include <chrono>
#include <cstdlib>
#include <iostream>
#include "scene_io.h"

using namespace mitsuba;

static double now() {
	return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

/* ------------------------------------------------------------------------------------ */
/* A film that keeps the raw {spec, alpha, weight} accumulators (exrfilm.cpp:37-44 keeps  */
/* the same triple per pixel) so they can be dumped without an image codec.               */
class RawFilm : public Film {
public:
	RawFilm(const Properties &props) : Film(props) {
		m_acc.assign((size_t) m_cropSize.x * m_cropSize.y * 5, 0.0f);
	}
	void putImageBlock(const ImageBlock *block) {
		const int b = block->getBorder();
		const Vector2i full = block->getFullSize();
		for (int y = 0; y < full.y; ++y) {
			const int iy = block->getOffset().y - b - m_cropOffset.y + y;
			if (iy < 0 || iy >= m_cropSize.y) continue;
			for (int x = 0; x < full.x; ++x) {
				const int ix = block->getOffset().x - b - m_cropOffset.x + x;
				if (ix < 0 || ix >= m_cropSize.x) continue;
				const size_t e = (size_t) y * full.x + x;
				float *px = &m_acc[((size_t) iy * m_cropSize.x + ix) * 5];
				Float r, g, bl;
				block->getPixel(e).toLinearRGB(r, g, bl);
				px[0] += r; px[1] += g; px[2] += bl;
				px[3] += block->getAlpha(e);
				px[4] += block->getWeight(e);
			}
		}
	}
	Spectrum getValue(int x, int y) {
		const float *px = &m_acc[((size_t) (y - m_cropOffset.y) * m_cropSize.x + (x - m_cropOffset.x)) * 5];
		Spectrum s; s.fromLinearRGB(px[0], px[1], px[2]);
		return px[4] != 0 ? s / px[4] : Spectrum(0.0f);
	}
	void clear() { std::fill(m_acc.begin(), m_acc.end(), 0.0f); }
	void fromBitmap(const Bitmap *bitmap) {
		const float *d = bitmap->getFloatData();
		for (size_t i = 0; i < (size_t) m_cropSize.x * m_cropSize.y; ++i) {
			m_acc[i*5+0] = d[i*4+0]; m_acc[i*5+1] = d[i*4+1]; m_acc[i*5+2] = d[i*4+2];
			m_acc[i*5+3] = d[i*4+3]; m_acc[i*5+4] = 1.0f;
		}
	}
	void toBitmap(Bitmap *) const { }
	void develop(const fs::path &) { }
	bool destinationExists(const fs::path &) const { return false; }
	const std::vector<float> &raw() const { return m_acc; }
	MTS_DECLARE_CLASS()
protected:
	virtual ~RawFilm() { }
	std::vector<float> m_acc;
};
MTS_IMPLEMENT_CLASS(RawFilm, false, Film)

/* ------------------------------------------------------------------------------------ */
struct Built {
	ref<Scene> scene;
	ref<RawFilm> film;
	std::vector<ref<BSDF> > bsdfs;       /* one instance per material id (for the bsdf command) */
	double buildSeconds;
};

static ConfigurableObject *create(const Class *cls, const Properties &props) {
	return PluginManager::getInstance()->createObject(cls, props);
}

static Spectrum rgb(const float *v) { Spectrum s; s.fromLinearRGB(v[0], v[1], v[2]); return s; }

static const std::vector<sio::Texture> *g_textures = NULL;   /* texture table of the scene being built */

/* <texture type="checkerboard"> (src/textures/checkerboard.cpp:30-33, src/librender/texture.cpp:43-56) */
static ref<Texture> makeTexture(const sio::Texture &t) {
	if (t.type == 1 && !t.exrFile.empty()) {      /* <texture type="exrtexture"> (exrtexture.cpp:34-44) */
		Properties p("exrtexture");
		p.setString("filename", t.exrFile);
		p.setFloat("uoffset", t.p[6]); p.setFloat("voffset", t.p[7]);
		p.setFloat("uscale", t.p[8]); p.setFloat("vscale", t.p[9]);
		ref<Texture> tex = static_cast<Texture *>(create(MTS_CLASS(Texture), p));
		tex->configure();
		return tex;
	}
	if (t.type == 1) {
		/* <texture type="ldrtexture"> reading a 24-bit BMP file written here from the scene file's pixels
		 * (src/textures/ldrtexture.cpp:37-84) */
		static int serial = 0;
		char path[256];
		snprintf(path, sizeof(path), "/tmp/mts_oracle_tex_%d_%d.bmp", (int) getpid(), serial++);
		const uint32_t line = t.width * 3, padded = (line + 3) & ~3u, dataSize = padded * t.height;
		std::vector<uint8_t> file(54 + dataSize, 0);
		auto put32 = [&](size_t at, uint32_t v) { for (int k = 0; k < 4; ++k) file[at + k] = (uint8_t) (v >> (8 * k)); };
		file[0] = 'B'; file[1] = 'M'; put32(2, (uint32_t) file.size()); put32(10, 54);
		put32(14, 40); put32(18, t.width); put32(22, t.height); file[26] = 1; file[28] = 24; put32(34, dataSize);
		for (uint32_t y = 0; y < t.height; ++y)               /* bottom-up rows, BGR */
			for (uint32_t x = 0; x < t.width; ++x)
				for (int c = 0; c < 3; ++c)
					file[54 + (size_t) (t.height - 1 - y) * padded + 3 * x + c] = t.pixels[((size_t) y * t.width + x) * 3 + (2 - c)];
		FILE *f = fopen(path, "wb");
		if (!f || fwrite(file.data(), 1, file.size(), f) != file.size()) SLog(EError, "cannot write %s", path);
		fclose(f);
		static const char *filters[3] = { "ewa", "trilinear", "none" };
		static const char *wraps[4] = { "repeat", "black", "white", "clamp" };
		Properties p("ldrtexture");
		p.setString("filename", path);
		p.setFloat("gamma", t.gamma);
		p.setString("filterType", filters[(int) t.p[0]]);
		p.setString("wrapMode", wraps[(int) t.p[1]]);
		p.setFloat("maxAnisotropy", t.p[2]);
		p.setFloat("uoffset", t.p[6]); p.setFloat("voffset", t.p[7]);
		p.setFloat("uscale", t.p[8]); p.setFloat("vscale", t.p[9]);
		ref<Texture> tex = static_cast<Texture *>(create(MTS_CLASS(Texture), p));
		tex->configure();
		/* LDRTexture::serialize re-reads the file (ldrtexture.cpp:218-227; the gpupath plugin goes through it), so the
		 * file stays until the process ends */
		static std::vector<std::string> files;
		if (files.empty()) atexit([]() { for (size_t k = 0; k < files.size(); ++k) remove(files[k].c_str()); });
		files.push_back(path);
		return tex;
	}
	if (t.type != 0 && t.type != 2) SLog(EError, "unknown texture type %u", t.type);
	Properties p(t.type == 2 ? "gridtexture" : "checkerboard");
	if (t.type == 2) p.setFloat("lineWidth", t.p[10]);
	p.setSpectrum("brightColor", rgb(t.p));
	p.setSpectrum("darkColor", rgb(t.p + 3));
	p.setFloat("uoffset", t.p[6]); p.setFloat("voffset", t.p[7]);
	p.setFloat("uscale", t.p[8]); p.setFloat("vscale", t.p[9]);
	ref<Texture> tex = static_cast<Texture *>(create(MTS_CLASS(Texture), p));
	tex->configure();
	return tex;
}

static ref<BSDF> makeBSDF(const sio::Material &mIn) {
	if (mIn.type & 0x200u) {
		/* <bsdf type="mask"> <spectrum name="opacity"/> <bsdf .../> </bsdf>  (mask.cpp:30-33, 162-170) */
		sio::Material inner = mIn; inner.type &= ~0x200u; inner.tex[2] = 0;
		ref<BSDF> nested = makeBSDF(inner);
		Properties mp("mask");
		mp.setSpectrum("opacity", rgb(mIn.p + 13));
		ref<BSDF> outer = static_cast<BSDF *>(create(MTS_CLASS(BSDF), mp));
		if (mIn.tex[2]) {
			if (!g_textures || mIn.tex[2] > g_textures->size()) SLog(EError, "bad texture reference");
			ref<Texture> tex = makeTexture((*g_textures)[mIn.tex[2] - 1]);
			outer->addChild("opacity", tex);
			tex->setParent(outer);
		}
		outer->addChild("", nested);
		nested->setParent(outer);
		outer->configure();
		return outer;
	}
	if (mIn.type & 0x100u) {
		/* <bsdf type="twosided"> <bsdf type="..."/> </bsdf>  (twosided.cpp:30-35) */
		sio::Material inner = mIn; inner.type &= 0xFFu;
		ref<BSDF> nested = makeBSDF(inner);
		ref<BSDF> outer = static_cast<BSDF *>(create(MTS_CLASS(BSDF), Properties("twosided")));
		outer->addChild("", nested);
		nested->setParent(outer);
		outer->configure();
		return outer;
	}
	const sio::Material &m = mIn;
	Properties p;
	switch (m.type) {
		case sio::MAT_LAMBERTIAN:
			p = Properties("lambertian");
			p.setSpectrum("reflectance", rgb(m.p));
			break;
		case sio::MAT_MICROFACET:
			p = Properties("microfacet");
			p.setSpectrum("diffuseReflectance", rgb(m.p));
			p.setSpectrum("specularReflectance", rgb(m.p + 3));
			p.setFloat("diffuseAmount", m.p[6]);
			p.setFloat("specularAmount", m.p[7]);
			p.setFloat("alphaB", m.p[8]);
			p.setFloat("intIOR", m.p[9]);
			p.setFloat("extIOR", m.p[10]);
			break;
		case sio::MAT_MIRROR:
			p = Properties("mirror");
			p.setSpectrum("specularReflectance", rgb(m.p));
			break;
		case sio::MAT_ROUGHMETAL:
			p = Properties("roughmetal");
			p.setSpectrum("specularReflectance", rgb(m.p));
			p.setFloat("alphaB", m.p[3]);
			p.setSpectrum("ior", rgb(m.p + 4));
			p.setSpectrum("k", rgb(m.p + 7));
			break;
		case sio::MAT_DIELECTRIC:
			p = Properties("dielectric");
			p.setSpectrum("specularReflectance", rgb(m.p));
			p.setSpectrum("specularTransmittance", rgb(m.p + 3));
			p.setFloat("intIOR", m.p[6]);
			p.setFloat("extIOR", m.p[7]);
			break;
		case 5: /* phong: the file holds post-configure values (phong.cpp:79-99) */
			p = Properties("phong");
			p.setSpectrum("diffuseReflectance", rgb(m.p));
			p.setSpectrum("specularReflectance", rgb(m.p + 3));
			p.setFloat("diffuseAmount", m.p[6]);
			p.setFloat("specularAmount", m.p[7]);
			p.setFloat("exponent", m.p[8]);
			p.setFloat("specularSamplingWeight", m.p[9]);
			p.setBoolean("verifyEnergyConservation", false);
			break;
		case 6: { /* roughglass: p[6] is the stored alpha-texture value; the phong distribution derives it from
		             the user-facing roughness kept in p[10] (roughglass.cpp:132-138) */
			static const char *distr[3] = { "beckmann", "phong", "ggx" };
			const int d = (int) m.p[9];
			p = Properties("roughglass");
			p.setSpectrum("specularReflectance", rgb(m.p));
			p.setSpectrum("specularTransmittance", rgb(m.p + 3));
			p.setFloat("alpha", d == 1 ? m.p[10] : m.p[6]);
			p.setFloat("intIOR", m.p[7]);
			p.setFloat("extIOR", m.p[8]);
			p.setString("distribution", distr[d < 0 || d > 2 ? 0 : d]);
			break;
		}
		case 7: { /* ward: the file holds post-configure values (ward.cpp:120-139) */
			static const char *models[3] = { "ward", "ward-duer", "balanced" };
			const int mt = (int) m.p[12];
			p = Properties("ward");
			p.setSpectrum("diffuseReflectance", rgb(m.p));
			p.setSpectrum("specularReflectance", rgb(m.p + 3));
			p.setFloat("diffuseAmount", m.p[6]);
			p.setFloat("specularAmount", m.p[7]);
			p.setFloat("alphaX", m.p[8]);
			p.setFloat("alphaY", m.p[9]);
			p.setFloat("specularSamplingWeight", m.p[10]);
			p.setString("type", models[mt < 0 || mt > 2 ? 2 : mt]);
			p.setBoolean("verifyEnergyConservation", false);
			break;
		}
		case 8: /* difftrans (difftrans.cpp:64-73) */
			p = Properties("difftrans");
			p.setSpectrum("transmittance", rgb(m.p));
			break;
		default:
			SLog(EError, "unknown material type %u", m.type);
	}
	ref<BSDF> bsdf = static_cast<BSDF *>(create(MTS_CLASS(BSDF), p));
	/* textured slots: <ref name="reflectance" id="..."/> inside the <bsdf> element */
	static const char *slotNames[9][2] = { { "reflectance", NULL }, { "diffuseReflectance", "specularReflectance" }, { NULL, NULL },
		{ "specularReflectance", NULL }, { NULL, NULL }, { "diffuseReflectance", "specularReflectance" },
		{ "specularReflectance", "specularTransmittance" }, { "diffuseReflectance", "specularReflectance" }, { "transmittance", NULL } };
	for (int k = 0; k < 2; ++k) {
		if (!m.tex[k]) continue;
		if (!g_textures || m.tex[k] > g_textures->size() || !slotNames[m.type][k]) SLog(EError, "bad texture reference");
		ref<Texture> tex = makeTexture((*g_textures)[m.tex[k] - 1]);
		bsdf->addChild(slotNames[m.type][k], tex);
		tex->setParent(bsdf);
	}
	bsdf->configure();
	return bsdf;
}

static Built buildScene(const sio::Scene &s, const std::string &integratorOverride, bool withGeometry = true) {
	Built out;
	g_textures = &s.textures;
	Properties sceneProps("scene");
	sceneProps.setBoolean("importanceSampleLuminaires", s.importanceSampleLuminaires != 0);     /* scene.cpp:51 */
	ref<Scene> scene = new Scene(sceneProps);

	/* integrator */
	{
		std::string type = integratorOverride.empty() ? (s.mode == 0 ? "path" : "direct") : integratorOverride;
		Properties p(type);
		if (integratorOverride.empty()) {
			if (s.mode == 0) {
				p.setInteger("maxDepth", s.maxDepth);
				p.setInteger("rrDepth", s.rrDepth);
				p.setBoolean("strictNormals", s.strictNormals != 0);
			} else {
				p.setInteger("luminaireSamples", (int) s.nLumSamples);
				p.setInteger("bsdfSamples", (int) s.nBsdfSamples);
			}
		} else {
			p.setString("mode", s.mode == 0 ? "path" : "direct");
			p.setInteger("maxDepth", s.maxDepth);
			p.setInteger("rrDepth", s.rrDepth);
			p.setBoolean("strictNormals", s.strictNormals != 0);
			p.setInteger("luminaireSamples", (int) s.nLumSamples);
			p.setInteger("bsdfSamples", (int) s.nBsdfSamples);
			if (getenv("MTS_ORACLE_GPUS")) p.setInteger("gpus", atoi(getenv("MTS_ORACLE_GPUS")));
			if (getenv("MTS_ORACLE_REFRESH")) p.setFloat("refreshInterval", (Float) atof(getenv("MTS_ORACLE_REFRESH")));
		}
		ref<Integrator> integrator = static_cast<Integrator *>(create(MTS_CLASS(Integrator), p));
		scene->addChild("", integrator);
		integrator->setParent(scene);
		integrator->configure();
	}

	/* film + filter, sampler, camera */
	{
		Properties fp("rawfilm");
		fp.setInteger("width", (int) s.width); fp.setInteger("height", (int) s.height);
		fp.setInteger("cropOffsetX", (int) s.cropX); fp.setInteger("cropOffsetY", (int) s.cropY);
		fp.setInteger("cropWidth", (int) s.cropW); fp.setInteger("cropHeight", (int) s.cropH);
		out.film = new RawFilm(fp);
		ref<ReconstructionFilter> filter = static_cast<ReconstructionFilter *>(
			create(MTS_CLASS(ReconstructionFilter), Properties(s.filterType == 0 ? "box" : s.filterType == 2 ? "mitchell" : s.filterType == 3 ? "catmullrom" : s.filterType == 4 ? "wsinc" : "gaussian")));
		filter->configure();
		out.film->addChild("", filter);
		filter->setParent(out.film);
		out.film->configure();

		Properties sp(s.samplerType == 1 ? "ldsampler" : s.samplerType == 2 ? "stratified" : s.samplerType == 3 ? "halton" : s.samplerType == 4 ? "hammersley" : "independent");
		sp.setInteger("sampleCount", (int) s.spp);
		if (s.samplerType == 1 || s.samplerType == 2) sp.setInteger("depth", (int) s.samplerDepth);
		ref<Sampler> sampler = static_cast<Sampler *>(create(MTS_CLASS(Sampler), sp));
		sampler->configure();

		Properties cp((s.mapSmallerSide & 2) ? "orthographic" : "perspective");     /* bit 1 of the flag word: orthographic camera */
		cp.setTransform("toWorld", Transform(Matrix4x4(s.cameraToWorld)));
		cp.setFloat("fov", s.fov);
		cp.setFloat("nearClip", s.nearClip);
		cp.setFloat("farClip", s.farClip);
		cp.setFloat("apertureRadius", s.apertureRadius);
		cp.setFloat("focusDepth", s.focusDepth);
		cp.setFloat("shutterOpen", s.shutterOpen);                 /* camera.cpp:28-35 */
		cp.setFloat("shutterClose", s.shutterClose);
		cp.setBoolean("mapSmallerSide", (s.mapSmallerSide & 1) != 0);
		ref<Camera> camera = static_cast<Camera *>(create(MTS_CLASS(Camera), cp));
		camera->addChild("", sampler);  sampler->setParent(camera);
		camera->addChild("", out.film); out.film->setParent(camera);
		camera->configure();
		scene->addChild("", camera);
		camera->setParent(scene);
	}

	for (size_t i = 0; i < s.materials.size(); ++i)
		out.bsdfs.push_back(makeBSDF(s.materials[i]));

	/* luminaires that are scene children (point.cpp, spot.cpp, directional.cpp, constant.cpp) */
	for (size_t i = 0; i < s.emitters.size(); ++i) {
		const sio::Emitter &e = s.emitters[i];
		if (e.type == 0) continue;
		static const char *names[7] = { "area", "point", "spot", "directional", "constant", "envmap", "collimated" };
		if (e.type > 6) SLog(EError, "unknown emitter type %u", e.type);
		Properties lp(names[e.type]);
		if (e.type == 5) {       /* <luminaire type="envmap"> (envmap.cpp:40-51) */
			lp.setString("filename", e.envFile);
			lp.setFloat("intensityScale", e.p[4]);
			lp.setTransform("toWorld", Transform(Matrix4x4(const_cast<float *>(e.aux))));
		} else
		lp.setSpectrum("intensity", rgb(e.intensity));
		lp.setFloat("samplingWeight", e.samplingWeight);
		if (e.type == 1 || e.type == 2 || e.type == 6)
			lp.setTransform("toWorld", Transform(Matrix4x4(const_cast<float *>(e.aux))));
		if (e.type == 6) lp.setFloat("radius", e.p[0]);
		if (e.type == 2) { lp.setFloat("beamWidth", e.p[8]); lp.setFloat("cutoffAngle", e.p[9]); }
		if (e.type == 3) lp.setPoint("direction", Point(e.aux[0], e.aux[1], e.aux[2]));   /* getVector reads the point-typed entry */
		ref<Luminaire> lum = static_cast<Luminaire *>(create(MTS_CLASS(Luminaire), lp));
		lum->configure();
		scene->addChild("", lum);
		lum->setParent(scene);
	}

	/* <shape type="shapegroup" id="..."> with its meshes as children (shapegroup.cpp:57-82); built on configure() */
	std::vector<ref<Shape> > groups;
	for (size_t g = 0; withGeometry && g < s.groups.size(); ++g) {
		Properties gp("shapegroup");
		gp.setID(formatString("group%i", (int) g));
		ref<Shape> group = static_cast<Shape *>(create(MTS_CLASS(Shape), gp));
		for (size_t i = 0; i < s.groups[g].size(); ++i) {
			const sio::Shape &sh = s.groups[g][i];
			ref<TriMesh> mesh = new TriMesh(formatString("group%i_%i", (int) g, (int) i), sh.nTri, sh.nVert,
				sh.hasNormals != 0, sh.hasUv != 0, false, false, sh.hasNormals == 0);
			if (sh.hasUv) memcpy(mesh->getVertexTexcoords(), sh.uv.data(), sh.uv.size() * 4);
			memcpy(mesh->getVertexPositions(), sh.pos.data(), sh.pos.size() * 4);
			if (sh.hasNormals) memcpy(mesh->getVertexNormals(), sh.nrm.data(), sh.nrm.size() * 4);
			memcpy(mesh->getTriangles(), sh.tri.data(), sh.tri.size() * 4);
			if (sh.matId >= 0) {
				ref<BSDF> bsdf = makeBSDF(s.materials[sh.matId]);
				mesh->addChild("", bsdf);
				bsdf->setParent(mesh);
			}
			mesh->configure();
			group->addChild("", mesh);
			mesh->setParent(group);
		}
		group->configure();
		groups.push_back(group);
	}

	if (withGeometry) {
		for (size_t i = 0; i < s.shapes.size(); ++i) {
			const sio::Shape &sh = s.shapes[i];
			if (sh.kind == 1) {      /* <shape type="sphere"> with center + radius (sphere.cpp:44-47) */
				Properties sp("sphere");
				sp.setPoint("center", Point(sh.center[0], sh.center[1], sh.center[2]));
				sp.setFloat("radius", sh.radius);
				sp.setBoolean("inverted", sh.inverted != 0);
				ref<Shape> sphere = static_cast<Shape *>(create(MTS_CLASS(Shape), sp));
				if (sh.matId >= 0) {
					ref<BSDF> bsdf = makeBSDF(s.materials[sh.matId]);
					sphere->addChild("", bsdf);
					bsdf->setParent(sphere);
				}
				if (sh.emitId >= 0) {    /* an <luminaire type="area"> nested in the sphere */
					const sio::Emitter &e = s.emitters[sh.emitId];
					Properties lp("area");
					lp.setSpectrum("intensity", rgb(e.intensity));
					lp.setFloat("samplingWeight", e.samplingWeight);
					ref<Luminaire> lum = static_cast<Luminaire *>(create(MTS_CLASS(Luminaire), lp));
					sphere->addChild("", lum);
					lum->setParent(sphere);
					lum->configure();
				}
				sphere->configure();
				scene->addChild("", sphere);
				sphere->setParent(scene);
				continue;
			}
			if (sh.kind == 3) {      /* <shape type="cylinder"> <transform name="toWorld"/> </shape> (cylinder.cpp:52-58) */
				Properties cp("cylinder");
				cp.setTransform("toWorld", Transform(Matrix4x4(const_cast<float *>(sh.toWorld))));
				ref<Shape> cyl = static_cast<Shape *>(create(MTS_CLASS(Shape), cp));
				if (sh.matId >= 0) {
					ref<BSDF> bsdf = makeBSDF(s.materials[sh.matId]);
					cyl->addChild("", bsdf);
					bsdf->setParent(cyl);
				}
				cyl->configure();
				scene->addChild("", cyl);
				cyl->setParent(scene);
				continue;
			}
			if (sh.kind == 2) {      /* <shape type="instance"> <ref id="group"/> <transform name="toWorld"/> </shape> (instance.cpp:22-26) */
				Properties ip("instance");
				ip.setTransform("toWorld", Transform(Matrix4x4(const_cast<float *>(sh.toWorld))));
				ref<Shape> inst = static_cast<Shape *>(create(MTS_CLASS(Shape), ip));
				inst->addChild("", groups[sh.groupId]);
				inst->configure();
				scene->addChild("", inst);
				inst->setParent(scene);
				continue;
			}
			ref<TriMesh> mesh = new TriMesh(formatString("shape%i", (int) i), sh.nTri, sh.nVert,
				sh.hasNormals != 0, sh.hasUv != 0, false, false, sh.hasNormals == 0 /* faceNormals */);
			if (sh.hasUv)
				memcpy(mesh->getVertexTexcoords(), sh.uv.data(), sh.uv.size() * 4);
			memcpy(mesh->getVertexPositions(), sh.pos.data(), sh.pos.size() * 4);
			if (sh.hasNormals)
				memcpy(mesh->getVertexNormals(), sh.nrm.data(), sh.nrm.size() * 4);
			memcpy(mesh->getTriangles(), sh.tri.data(), sh.tri.size() * 4);
			if (sh.matId >= 0) {
				/* a fresh instance per shape, like one <bsdf> element nested in each <shape> */
				ref<BSDF> bsdf = makeBSDF(s.materials[sh.matId]);
				mesh->addChild("", bsdf);
				bsdf->setParent(mesh);
			}
			if (sh.emitId >= 0) {
				const sio::Emitter &e = s.emitters[sh.emitId];
				Properties lp("area");
				lp.setSpectrum("intensity", rgb(e.intensity));
				lp.setFloat("samplingWeight", e.samplingWeight);
				ref<Luminaire> lum = static_cast<Luminaire *>(create(MTS_CLASS(Luminaire), lp));
				mesh->addChild("", lum);
				lum->setParent(mesh);
				lum->configure();
			}
			mesh->configure();
			scene->addChild("", mesh);
			mesh->setParent(scene);
		}
	}
	scene->configure();
	double t0 = now();
	if (withGeometry)
		scene->initialize();
	out.buildSeconds = now() - t0;
	out.scene = scene;
	return out;
}

/* ------------------------------------------------------------------------------------ */
/* Worker threads are mitsuba Threads so Thread::getID() (used by StatsCounter) is valid. */
class RangeThread : public Thread {
public:
	typedef void (*Fn)(void *ctx, size_t begin, size_t end);
	RangeThread(const std::string &name, Fn fn, void *ctx, size_t b, size_t e)
		: Thread(name), m_fn(fn), m_ctx(ctx), m_b(b), m_e(e) { }
	void run() { m_fn(m_ctx, m_b, m_e); }
	MTS_DECLARE_CLASS()
protected:
	virtual ~RangeThread() { }
	Fn m_fn; void *m_ctx; size_t m_b, m_e;
};
MTS_IMPLEMENT_CLASS(RangeThread, false, Thread)

static void parallelFor(size_t n, int threads, RangeThread::Fn fn, void *ctx) {
	if (threads <= 1) { fn(ctx, 0, n); return; }
	std::vector<ref<RangeThread> > pool;
	for (int i = 0; i < threads; ++i) {
		size_t b = n * i / threads, e = n * (i + 1) / threads;
		pool.push_back(new RangeThread(formatString("trc%i", i), fn, ctx, b, e));
		pool.back()->start();
	}
	for (size_t i = 0; i < pool.size(); ++i)
		pool[i]->join();
}

/* ShapeKDTree befriends a class called mitsuba::Instance (skdtree.h:70; the real one lives
 * in the out-of-scope 'instance' shape plugin). Defining our own class of that name gives the
 * harness legitimate access to the protected traversal routine and the IntersectionCache, so
 * it can report (shapeIndex, primIndex, u, v) — which the public Intersection record drops. */
MTS_NAMESPACE_BEGIN
class Instance {
public:
	/* Same pre-amble as ShapeKDTree::rayIntersect(ray, its) (skdtree.cpp:108-124), then the
	 * reference's own rayIntersectHavran<false>. */
	static bool closest(const ShapeKDTree *kd, const Ray &ray, Float &t,
			uint32_t &shape, uint32_t &prim, Float &u, Float &v) {
		uint8_t temp[MTS_KD_INTERSECTION_TEMP];
		t = std::numeric_limits<Float>::infinity();
		Float mint, maxt;
		if (kd->m_aabb.rayIntersect(ray, mint, maxt)) {
			Float rayMinT = ray.mint;
			if (rayMinT == Epsilon)
				rayMinT *= std::max(std::max(std::max(std::abs(ray.o.x),
					std::abs(ray.o.y)), std::abs(ray.o.z)), Epsilon);
			if (rayMinT > mint) mint = rayMinT;
			if (ray.maxt < maxt) maxt = ray.maxt;
			if (maxt > mint && kd->rayIntersectHavran<false>(ray, mint, maxt, t, temp)) {
				const ShapeKDTree::IntersectionCache *c =
					reinterpret_cast<const ShapeKDTree::IntersectionCache *>(temp);
				shape = c->shapeIndex; prim = c->primIndex; u = c->u; v = c->v;
				return true;
			}
		}
		return false;
	}
	/* The same query through the reference's own instrumented traversal (sahkdtree3.h:310-410): kd-tree nodes traversed
	 * and primitives intersected for this ray — the per-ray work the GPU side's counters are set against. */
	static bool closestStats(const ShapeKDTree *kd, const Ray &ray, Float &t, uint32_t &traversals, uint32_t &intersections) {
		uint8_t temp[MTS_KD_INTERSECTION_TEMP];
		t = std::numeric_limits<Float>::infinity();
		traversals = intersections = 0;
		Float mint, maxt;
		if (!kd->m_aabb.rayIntersect(ray, mint, maxt)) return false;
		Float rayMinT = ray.mint;
		if (rayMinT == Epsilon)
			rayMinT *= std::max(std::max(std::max(std::abs(ray.o.x), std::abs(ray.o.y)), std::abs(ray.o.z)), Epsilon);
		if (rayMinT > mint) mint = rayMinT;
		if (ray.maxt < maxt) maxt = ray.maxt;
		if (!(maxt > mint)) return false;
		boost::tuple<bool, uint32_t, uint32_t, uint64_t> st = kd->rayIntersectHavranCollectStatistics(ray, mint, maxt, t, temp);
		traversals = boost::get<1>(st); intersections = boost::get<2>(st);
		return boost::get<0>(st);
	}
	static const TriAccel *triAccel(const ShapeKDTree *kd) { return kd->m_triAccel; }
	static const AABB &tightAABB(const ShapeKDTree *kd) { return kd->m_tightAABB; }
};
MTS_NAMESPACE_END

struct TraceCtx {
	const Scene *scene; const float *rays; int mode;
	float *tuv; uint32_t *ids; float *full; uint32_t *occl;
	std::atomic<uint64_t> *kdTraversals, *kdIntersections;     /* mode 3 */
};

static void traceRange(void *c, size_t b, size_t e) {
	TraceCtx &ctx = *static_cast<TraceCtx *>(c);
	const ShapeKDTree *kd = ctx.scene->getKDTree();
	for (size_t i = b; i < e; ++i) {
		const float *r = ctx.rays + i * 8;
		/* same constructor the integrators use: dRcp = 1/d per component (ray.h:78-91) */
		Ray ray(Point(r[0], r[1], r[2]), Vector(r[4], r[5], r[6]), r[3], r[7], 0.0f);
		if (ctx.mode == 2) {
			ctx.occl[i] = kd->rayIntersect(ray) ? 1u : 0u;
			continue;
		}
		if (ctx.mode == 3) {
			Float t; uint32_t nt, ni;
			const bool hit = Instance::closestStats(kd, ray, t, nt, ni);
			ctx.tuv[i*3+0] = hit ? t : std::numeric_limits<float>::infinity(); ctx.tuv[i*3+1] = ctx.tuv[i*3+2] = 0;
			ctx.ids[i*2+0] = nt; ctx.ids[i*2+1] = ni;
			ctx.kdTraversals->fetch_add(nt, std::memory_order_relaxed); ctx.kdIntersections->fetch_add(ni, std::memory_order_relaxed);
			continue;
		}
		Float t, u, v; uint32_t shape, prim;
		if (Instance::closest(kd, ray, t, shape, prim, u, v)) {
			ctx.tuv[i*3+0] = t; ctx.tuv[i*3+1] = u; ctx.tuv[i*3+2] = v;
			ctx.ids[i*2+0] = shape; ctx.ids[i*2+1] = prim;
		} else {
			ctx.tuv[i*3+0] = std::numeric_limits<float>::infinity();
			ctx.tuv[i*3+1] = ctx.tuv[i*3+2] = 0;
			ctx.ids[i*2+0] = 0xFFFFFFFFu; ctx.ids[i*2+1] = 0xFFFFFFFFu;
		}
		if (ctx.mode == 1) {
			/* the public entry point, for the filled Intersection record (skdtree.h:352-432) */
			float *f = ctx.full + i * 12;
			Intersection its;
			if (kd->rayIntersect(ray, its)) {
				if (its.t != t) SLog(EError, "harness: public and friend closest-hit disagree");
				f[0] = its.p.x; f[1] = its.p.y; f[2] = its.p.z;
				f[3] = its.geoFrame.n.x; f[4] = its.geoFrame.n.y; f[5] = its.geoFrame.n.z;
				f[6] = its.shFrame.n.x; f[7] = its.shFrame.n.y; f[8] = its.shFrame.n.z;
				f[9] = its.wi.x; f[10] = its.wi.y; f[11] = its.wi.z;
			} else {
				for (int k = 0; k < 12; ++k) f[k] = 0;
			}
		}
	}
}

static int dispatch(int argc, char **argv);

/* ------------------------------------------------------------------------------------ */
static std::string statValue(const std::string &stats, const std::string &name) {
	size_t pos = stats.find(name);
	if (pos == std::string::npos) return "0";
	size_t colon = stats.find(':', pos);
	size_t eol = stats.find('\n', colon);
	std::string v = stats.substr(colon + 1, eol - colon - 1);
	size_t a = v.find_first_not_of(" \t");
	size_t b = v.find_last_not_of(" \t\r");
	return a == std::string::npos ? "0" : v.substr(a, b - a + 1);
}

int main(int argc, char **argv) {
	Class::staticInitialization();
	PluginManager::staticInitialization();
	Statistics::staticInitialization();
	Thread::staticInitialization();
	Logger::staticInitialization();
	Spectrum::staticInitialization();
	Scheduler::staticInitialization();
	int rv = 1;
	try {
		Thread::getThread()->getLogger()->setLogLevel(getenv("MTS_ORACLE_VERBOSE") ? EDebug : EWarn);
		/* plugins/<name>.so live next to this executable */
		fs::path exe = fs::complete(fs::path(argv[0]));
		char self[4096]; ssize_t n = readlink("/proc/self/exe", self, sizeof(self) - 1);
		if (n > 0) { self[n] = 0; exe = fs::path(std::string(self)); }
		Thread::getThread()->getFileResolver()->addPath(exe.parent_path());
		if (getenv("MTS_PLUGIN_PATH"))
			Thread::getThread()->getFileResolver()->addPath(fs::path(std::string(getenv("MTS_PLUGIN_PATH"))));
		rv = dispatch(argc, argv);
	} catch (const std::exception &e) {
		std::cerr << "mts_oracle: " << e.what() << std::endl;
		rv = 2;
	}
	fflush(stdout);
	/* Skip the static shutdown sequence: worker threads may still hold references. */
	_exit(rv);
}

/* ------------------------------------------------------------------------------------ */
#include "commands.inl"

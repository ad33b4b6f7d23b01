/*
 * gpupath — Mitsuba 0.2.1 integrator plugin that runs the `path` / `direct` hot path on a B200 through
 * libmtscuda.so (C ABI: include/mtsgpu.h). Drop-in: same plugin ABI (MTS_EXPORT_PLUGIN,
 * include/mitsuba/core/cobject.h:79-87), same scene XML, same Film/Sampler/Camera interfaces.
 *
 *   <integrator type="gpupath"> <string name="mode" value="path"/> <integer name="maxDepth" value="8"/> ... </integrator>
 *
 * It subclasses SampleIntegrator (include/mitsuba/render/integrator.h:281-398) and overrides render(),
 * bypassing BlockedRenderProcess (src/librender/integrator.cpp:87-121); the precedent for an integrator
 * that renders on a device and hands the film a finished image is VPLIntegrator::render
 * (src/integrators/vpl/vpl.cpp:99-198). There is no CPU fallback: Li() raises an error.
 */
#include <mitsuba/render/scene.h>
#include <mitsuba/render/renderjob.h>
#include <mitsuba/render/renderqueue.h>
#include <mitsuba/render/trimesh.h>
#include <mitsuba/render/imageblock.h>
#include <mitsuba/core/mstream.h>
#include <mitsuba/core/bitmap.h>
#include <mitsuba/core/statistics.h>
#include <mitsuba/core/timer.h>
#include <map>
#include <vector>
#include <algorithm>
#include "mtsgpu.h"
#include "instance.h"          /* src/shapes/instance.h (+ shapegroup.h): Instance::getShapeGroup, ShapeGroup::getKDTree */

MTS_NAMESPACE_BEGIN

static StatsCounter gpuNormalRays("GPU path tracer", "Normal rays traced");
static StatsCounter gpuShadowRays("GPU path tracer", "Shadow rays traced");

class GPUPathIntegrator : public SampleIntegrator {
public:
	GPUPathIntegrator(const Properties &props) : SampleIntegrator(props), m_cancel(0) {
		std::string mode = props.getString("mode", "path");
		if (mode == "path") m_mode = MTSGPU_MODE_PATH;
		else if (mode == "direct") m_mode = MTSGPU_MODE_DIRECT;
		else Log(EError, "gpupath: unknown mode \"%s\" (must be \"path\" or \"direct\")", mode.c_str());
		/* same property names and defaults as MonteCarloIntegrator (integrator.cpp:270-295) and
		   MIDirectIntegrator (direct.cpp:34-36) */
		m_maxDepth = props.getInteger("maxDepth", -1);
		m_rrDepth = props.getInteger("rrDepth", 10);
		m_strictNormals = props.getBoolean("strictNormals", false);
		m_luminaireSamples = props.getInteger("luminaireSamples", 1);
		m_bsdfSamples = props.getInteger("bsdfSamples", 1);
		m_device = props.getInteger("device", 0);
		/* number of GPUs to shard the sample range over (devices device .. device+gpus-1) */
		m_gpus = props.getInteger("gpus", 1);
		m_seed = props.getInteger("seed", 0);
		/* seconds between film previews (RenderQueue::signalRefresh) while the GPU renders; 0: after every wavefront batch */
		m_refreshInterval = props.getFloat("refreshInterval", 2.0f);
		AssertEx(m_rrDepth > 0, "rrDepth == 0 breaks the computation of alpha values!");
		Assert(m_luminaireSamples + m_bsdfSamples > 0);
	}

	GPUPathIntegrator(Stream *stream, InstanceManager *manager) : SampleIntegrator(stream, manager), m_cancel(0) {
		m_mode = stream->readInt(); m_maxDepth = stream->readInt(); m_rrDepth = stream->readInt();
		m_strictNormals = stream->readBool(); m_luminaireSamples = stream->readInt(); m_bsdfSamples = stream->readInt();
		m_device = stream->readInt(); m_seed = stream->readInt(); m_gpus = stream->readInt();
		m_refreshInterval = stream->readFloat();
	}

	void serialize(Stream *stream, InstanceManager *manager) const {
		SampleIntegrator::serialize(stream, manager);
		stream->writeInt(m_mode); stream->writeInt(m_maxDepth); stream->writeInt(m_rrDepth);
		stream->writeBool(m_strictNormals); stream->writeInt(m_luminaireSamples); stream->writeInt(m_bsdfSamples);
		stream->writeInt(m_device); stream->writeInt(m_seed); stream->writeInt(m_gpus);
		stream->writeFloat(m_refreshInterval);
	}

	Spectrum Li(const RayDifferential &, RadianceQueryRecord &) const {
		Log(EError, "gpupath has no CPU path: Li() must not be called (use the 'path' or 'direct' plugin on the CPU)");
		return Spectrum(0.0f);
	}

	void cancel() { m_cancel = 1; }

	bool render(Scene *scene, RenderQueue *queue, const RenderJob *job, int, int, int) {
		m_cancel = 0;
		std::vector<mtsgpu_ctx *> ctxs;
		bool ok = false;
		try {
			for (int g = 0; g < std::max(1, m_gpus); ++g) {
				mtsgpu_ctx *ctx = NULL;
				if (mtsgpu_create(m_device + g, &ctx) != MTSGPU_OK)
					Log(EError, "gpupath: GPU %i: %s", m_device + g, mtsgpu_last_error(NULL));
				ctxs.push_back(ctx);
			}
			ok = renderImpl(ctxs, scene, queue, job);
		} catch (...) {
			for (size_t i = 0; i < ctxs.size(); ++i) mtsgpu_destroy(ctxs[i]);
			throw;
		}
		for (size_t i = 0; i < ctxs.size(); ++i) mtsgpu_destroy(ctxs[i]);
		return ok;
	}

	MTS_DECLARE_CLASS()
private:
	void check(mtsgpu_ctx *ctx, int rc) const {
		/* Log(EError) throws std::runtime_error (src/libcore/logger.cpp:137); RenderJob::run catches it and
		   marks the job cancelled (src/librender/renderjob.cpp:126-130) */
		if (rc != MTSGPU_OK && rc != MTSGPU_ERR_CANCELLED)
			Log(EError, "gpupath: %s", mtsgpu_last_error(ctx));
	}

	/* ---- BSDF parameters: no getters exist, so decode the object's own wire format ---------------------- */
	static void put(float *dst, const Spectrum &v) { Float r, g, b; v.toLinearRGB(r, g, b); dst[0] = r; dst[1] = g; dst[2] = b; }

	/* Reading state for one serialised BSDF: the stream, the highest instance id met so far (InstanceManager::serialize
	 * writes id + class + payload the first time it meets an object and only the id afterwards,
	 * serialization.cpp:73-87), the textures collected for the C ABI, and which id became which slot value. */
	struct TexValue { float rgb[3]; uint32_t texture; };
	/* pixels of a bitmap texture, decoded by the host's own Bitmap codecs, waiting for mtsgpu_upload_bitmap */
	struct BitmapUpload { int texture; ref<Bitmap> bitmap; float gamma; };
	struct Decoder {
		Stream *s; unsigned int maxId;
		std::vector<mtsgpu_texture> *textures;
		std::vector<BitmapUpload> *bitmaps;
		std::map<unsigned int, TexValue> seen;
	};

	/* ConfigurableObject::serialize writes the parent (properties.cpp:358-363): nothing, a back-reference, or — for a
	 * texture shared with a BSDF that is not part of this stream — that BSDF inline, which is read and dropped. */
	static void readParent(Decoder &d) {
		const unsigned int id = d.s->readUInt();
		if (id == 0 || id <= d.maxId) return;
		d.maxId = id;
		const std::string cls = d.s->readString();
		decodePayload(cls, d);
	}

	/* One spectrum slot: a ConstantSpectrumTexture (texture.cpp:84-93), a Checkerboard (checkerboard.cpp:35-45) or an
	 * LDRTexture (ldrtexture.cpp:209-228) */
	static void readTextureSlot(Decoder &d, float *rgb, uint32_t &texture) {
		Stream *s = d.s;
		const unsigned int id = s->readUInt();
		if (id == 0)
			SLog(EError, "gpupath: BSDF without a texture");
		TexValue v; memset(&v, 0, sizeof(v));
		if (id <= d.maxId) {
			if (!d.seen.count(id)) SLog(EError, "gpupath: unexpected instance reference in a BSDF");
			v = d.seen[id];
		} else {
			d.maxId = id;
			const std::string cls = s->readString();
			readParent(d);
			if (cls == "ConstantSpectrumTexture") {
				put(v.rgb, Spectrum(s));
			} else if (cls == "Checkerboard") {
				mtsgpu_texture t; memset(&t, 0, sizeof(t));
				t.type = MTSGPU_TEXTURE_CHECKERBOARD;
				Point2 uvOffset(s); Vector2 uvScale(s);              /* Texture2D::serialize (texture.cpp:67-71) */
				put(t.p, Spectrum(s)); put(t.p + 3, Spectrum(s));    /* bright, dark */
				t.p[6] = uvOffset.x; t.p[7] = uvOffset.y; t.p[8] = uvScale.x; t.p[9] = uvScale.y;
				d.textures->push_back(t);
				v.texture = (uint32_t) d.textures->size();
			} else if (cls == "GridTexture") {                          /* gridtexture.cpp:43-48 */
				mtsgpu_texture t; memset(&t, 0, sizeof(t));
				t.type = MTSGPU_TEXTURE_GRID;
				Point2 uvOffset(s); Vector2 uvScale(s);
				put(t.p, Spectrum(s)); put(t.p + 3, Spectrum(s));
				t.p[10] = s->readFloat();
				t.p[6] = uvOffset.x; t.p[7] = uvOffset.y; t.p[8] = uvScale.x; t.p[9] = uvScale.y;
				d.textures->push_back(t);
				v.texture = (uint32_t) d.textures->size();
			} else if (cls == "LDRTexture") {
				/* ldrtexture.cpp:209-228: Texture2D, file name, gamma, codec, filter, wrap mode, max. anisotropy, then the
				 * image file itself — decoded here by the host's Bitmap class, exactly as LDRTexture's own
				 * unserialising constructor does (ldrtexture.cpp:86-100) */
				mtsgpu_texture t; memset(&t, 0, sizeof(t));
				t.type = MTSGPU_TEXTURE_BITMAP;
				Point2 uvOffset(s); Vector2 uvScale(s);
				s->readString();
				BitmapUpload up;
				up.gamma = s->readFloat();
				const Bitmap::EFileFormat format = static_cast<Bitmap::EFileFormat>(s->readInt());
				t.p[0] = (float) s->readInt(); t.p[1] = (float) s->readUInt(); t.p[2] = s->readFloat();
				t.p[6] = uvOffset.x; t.p[7] = uvOffset.y; t.p[8] = uvScale.x; t.p[9] = uvScale.y;
				const uint32_t size = s->readUInt();
				ref<MemoryStream> file = new MemoryStream(size);
				s->copyTo(file, size);
				file->setPos(0);
				up.bitmap = new Bitmap(format, file);
				d.textures->push_back(t);
				v.texture = (uint32_t) d.textures->size();
				up.texture = (int) d.textures->size() - 1;
				d.bitmaps->push_back(up);
			} else if (cls == "EXRTexture") {
				/* exrtexture.cpp:61-67: Texture2D, file name, then the OpenEXR file; MIPMap::fromBitmap's defaults
				 * (EWA, repeat, maxAnisotropy 8) */
				mtsgpu_texture t; memset(&t, 0, sizeof(t));
				t.type = MTSGPU_TEXTURE_BITMAP;
				Point2 uvOffset(s); Vector2 uvScale(s);
				s->readString();
				t.p[0] = 0; t.p[1] = 0; t.p[2] = 8.0f;
				t.p[6] = uvOffset.x; t.p[7] = uvOffset.y; t.p[8] = uvScale.x; t.p[9] = uvScale.y;
				const size_t size = s->readSize();
				ref<MemoryStream> file = new MemoryStream(size);
				s->copyTo(file, size);
				file->setPos(0);
				BitmapUpload up;
				up.gamma = 1.0f;
				up.bitmap = new Bitmap(Bitmap::EEXR, file);         /* 128 bpp: RGBA floats */
				d.textures->push_back(t);
				v.texture = (uint32_t) d.textures->size();
				up.texture = (int) d.textures->size() - 1;
				d.bitmaps->push_back(up);
			} else {
				SLog(EError, "gpupath: texture class \"%s\" is outside the supported hot path (constant spectra, checkerboard, ldrtexture, exrtexture)", cls.c_str());
			}
			d.seen[id] = v;
		}
		memcpy(rgb, v.rgb, sizeof(v.rgb));
		texture = v.texture;
	}

	/* `s` is positioned at the serialised form of an object of class `cls` */
	static mtsgpu_material decodePayload(const std::string &cls, Decoder &d) {
		Stream *s = d.s;
		mtsgpu_material m; memset(&m, 0, sizeof(m));
		readParent(d);                     /* BSDF::setParent is a no-op (bsdf.cpp:55-57) => normally 0 */
		s->readString();                   /* BSDF::serialize: m_name (bsdf.cpp:50-53) */
		if (cls == "TwoSidedBRDF") {                  /* twosided.cpp:55-59: the nested BRDF instance */
			const unsigned int id = s->readUInt();
			if (id == 0 || id <= d.maxId)
				SLog(EError, "gpupath: TwoSidedBRDF without a nested BRDF of its own");
			d.maxId = id;
			const std::string nested = s->readString();
			m = decodePayload(nested, d);
			if ((m.type & MTSGPU_BSDF_TYPE_MASK) == MTSGPU_BSDF_DIELECTRIC || (m.type & MTSGPU_BSDF_TYPE_MASK) == MTSGPU_BSDF_ROUGHGLASS
					|| (m.type & MTSGPU_BSDF_TWOSIDED))
				SLog(EError, "gpupath: TwoSidedBRDF around \"%s\" is not supported", nested.c_str());
			m.type |= MTSGPU_BSDF_TWOSIDED;
		} else if (cls == "Mask") {                   /* mask.cpp:47-51: the opacity texture, then the nested BSDF instance */
			float opacity[3]; uint32_t opacityTexture = 0;
			readTextureSlot(d, opacity, opacityTexture);
			const unsigned int id = s->readUInt();
			if (id == 0 || id <= d.maxId)
				SLog(EError, "gpupath: Mask without a nested BSDF of its own");
			d.maxId = id;
			const std::string nested = s->readString();
			m = decodePayload(nested, d);
			if (m.type & MTSGPU_BSDF_MASKED)
				SLog(EError, "gpupath: a Mask inside a Mask is not supported");
			m.type |= MTSGPU_BSDF_MASKED;
			memcpy(m.p + 13, opacity, sizeof(opacity));
			m.texture[2] = opacityTexture;
		} else if (cls == "DiffuseTransmitter") {     /* difftrans.cpp:76-79: the transmittance texture */
			m.type = MTSGPU_BSDF_DIFFTRANS;
			readTextureSlot(d, m.p, m.texture[0]);
		} else if (cls == "Lambertian") {             /* lambertian.cpp:137-141 */
			m.type = MTSGPU_BSDF_LAMBERTIAN;
			readTextureSlot(d, m.p, m.texture[0]);
		} else if (cls == "Microfacet") {             /* microfacet.cpp:283-293 */
			m.type = MTSGPU_BSDF_MICROFACET;
			readTextureSlot(d, m.p, m.texture[0]); readTextureSlot(d, m.p + 3, m.texture[1]);
			m.p[8] = s->readFloat(); m.p[6] = s->readFloat(); m.p[7] = s->readFloat();   /* alphaB, kd, ks */
			m.p[9] = s->readFloat(); m.p[10] = s->readFloat();                            /* intIOR, extIOR */
		} else if (cls == "Mirror") {                 /* mirror.cpp:51-55 */
			m.type = MTSGPU_BSDF_MIRROR;
			put(m.p, Spectrum(s));
		} else if (cls == "RoughMetal") {             /* roughmetal.cpp:169-176 */
			m.type = MTSGPU_BSDF_ROUGHMETAL;
			readTextureSlot(d, m.p, m.texture[0]);
			m.p[3] = s->readFloat();
			put(m.p + 4, Spectrum(s)); put(m.p + 7, Spectrum(s));
		} else if (cls == "Phong") {                  /* phong.cpp:247-257 */
			m.type = MTSGPU_BSDF_PHONG;
			readTextureSlot(d, m.p, m.texture[0]); readTextureSlot(d, m.p + 3, m.texture[1]);
			m.p[8] = s->readFloat(); m.p[6] = s->readFloat(); m.p[7] = s->readFloat();   /* exponent, kd, ks */
			m.p[9] = s->readFloat(); m.p[10] = s->readFloat();                            /* sampling weights */
		} else if (cls == "Ward") {                   /* ward.cpp:299-311 */
			m.type = MTSGPU_BSDF_WARD;
			m.p[12] = (float) s->readUInt();
			readTextureSlot(d, m.p, m.texture[0]); readTextureSlot(d, m.p + 3, m.texture[1]);
			m.p[8] = s->readFloat(); m.p[9] = s->readFloat(); m.p[6] = s->readFloat(); m.p[7] = s->readFloat();   /* alphaX, alphaY, kd, ks */
			m.p[10] = s->readFloat(); m.p[11] = s->readFloat();                                                    /* sampling weights */
		} else if (cls == "RoughGlass") {             /* roughglass.cpp:735-744 */
			m.type = MTSGPU_BSDF_ROUGHGLASS;
			m.p[9] = (float) s->readInt();
			{
				const unsigned int id = s->readUInt();
				if (id == 0 || id <= d.maxId || s->readString() != "ConstantFloatTexture")
					SLog(EError, "gpupath: RoughGlass: only a constant alpha is inside the supported hot path");
				d.maxId = id;
				readParent(d);
				m.p[6] = s->readFloat();               /* ConstantFloatTexture::serialize (texture.cpp:100-103) */
			}
			readTextureSlot(d, m.p, m.texture[0]); readTextureSlot(d, m.p + 3, m.texture[1]);
			m.p[7] = s->readFloat(); m.p[8] = s->readFloat();
		} else if (cls == "Dielectric") {             /* dielectric.cpp:88-95 */
			m.type = MTSGPU_BSDF_DIELECTRIC;
			m.p[6] = s->readFloat(); m.p[7] = s->readFloat();
			uint32_t t0 = 0, t1 = 0;
			readTextureSlot(d, m.p, t0); readTextureSlot(d, m.p + 3, t1);
			if (t0 || t1) SLog(EError, "gpupath: textured dielectric is outside the supported hot path");
		} else {
			SLog(EError, "gpupath: BSDF class \"%s\" is outside the supported hot path "
				"(lambertian, microfacet, mirror, roughmetal, dielectric, phong, roughglass, ward, twosided)", cls.c_str());
		}
		return m;
	}

	/* Scene-level luminaires (point, spot, directional, constant): no getters either, so decode
	 * Luminaire::serialize (luminaire.cpp:66-74) + the subclass's fields. ConfigurableObject::serialize writes the
	 * parent object first (properties.cpp:358-363) — the whole scene here — so the parent link is cleared for the
	 * duration of the call; render() runs before any worker touches the scene. */
	struct EnvUpload { ref<Bitmap> bitmap; float l2w[16], w2l[16]; };
	static void decodeLuminaire(const Luminaire *lumConst, mtsgpu_emitter &e, EnvUpload &env) {
		Luminaire *lum = const_cast<Luminaire *>(lumConst);
		const std::string cls = lum->getClass()->getName();
		ConfigurableObject *parent = lum->getParent();
		ref<MemoryStream> s = new MemoryStream();
		ref<InstanceManager> mgr = new InstanceManager();
		lum->setParent(NULL);
		lum->serialize(s, mgr);
		lum->setParent(parent);
		s->setPos(0);
		if (s->readUInt() != 0) SLog(EError, "gpupath: unexpected luminaire parent");
		if (s->readUInt() != 0) SLog(EError, "gpupath: luminaires inside participating media are outside the supported hot path");
		s->readFloat(); s->readInt(); s->readBool();           /* samplingWeight, type, intersectable */
		Transform worldToLuminaire(s);
		s->readString();
		const Matrix4x4 &w2l = worldToLuminaire.getMatrix(), &l2w = worldToLuminaire.getInverseMatrix();
		if (cls == "PointLuminaire") {                          /* point.cpp:41-45 */
			e.type = MTSGPU_EMITTER_POINT;
			put(e.intensity, Spectrum(s));
			const Point pos = worldToLuminaire.inverse()(Point(0, 0, 0));
			e.p[0] = pos.x; e.p[1] = pos.y; e.p[2] = pos.z;
		} else if (cls == "SpotLuminaire") {                    /* spot.cpp:63-70 */
			e.type = MTSGPU_EMITTER_SPOT;
			if (s->readUInt() == 0 || s->readString() != "ConstantSpectrumTexture" || s->readUInt() != 0)
				SLog(EError, "gpupath: SpotLuminaire: only a constant projection texture is inside the supported hot path");
			Spectrum tex(s), intensity(s);
			Float r, g, b; tex.toLinearRGB(r, g, b);
			if (r != 1 || g != 1 || b != 1)
				SLog(EError, "gpupath: SpotLuminaire: projection textures are outside the supported hot path");
			put(e.intensity, intensity);
			const Point pos = worldToLuminaire.inverse()(Point(0, 0, 0));
			e.p[0] = pos.x; e.p[1] = pos.y; e.p[2] = pos.z;
			e.p[3] = w2l.m[2][0]; e.p[4] = w2l.m[2][1]; e.p[5] = w2l.m[2][2];
			e.p[6] = s->readFloat(); e.p[7] = s->readFloat();  /* beamWidth, cutoffAngle (radians) */
		} else if (cls == "DirectionalLuminaire") {             /* directional.cpp:56-62 */
			e.type = MTSGPU_EMITTER_DIRECTIONAL;
			Vector d(s);
			put(e.intensity, Spectrum(s));
			Point diskOrigin(s);
			e.p[0] = d.x; e.p[1] = d.y; e.p[2] = d.z; e.p[3] = s->readFloat();
		} else if (cls == "ConstantLuminaire") {                /* constant.cpp:41-46, after preprocess() */
			e.type = MTSGPU_EMITTER_CONSTANT;
			put(e.intensity, Spectrum(s));
			BSphere bs(s);
			e.p[0] = bs.center.x; e.p[1] = bs.center.y; e.p[2] = bs.center.z; e.p[3] = bs.radius;
		} else if (cls == "CollimatedBeamLuminaire") {          /* collimated.cpp:46-50 */
			e.type = MTSGPU_EMITTER_COLLIMATED;
			put(e.intensity, Spectrum(s));
			e.p[0] = s->readFloat();
			for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { env.l2w[4 * i + j] = l2w.m[i][j]; env.w2l[4 * i + j] = w2l.m[i][j]; }
		} else if (cls == "EnvMapLuminaire") {                  /* envmap.cpp:79-92, after preprocess() */
			e.type = MTSGPU_EMITTER_ENVMAP;
			e.intensity[0] = e.intensity[1] = e.intensity[2] = 1.0f;
			e.p[4] = s->readFloat();                            /* intensityScale */
			s->readString();
			BSphere bs(s);
			e.p[0] = bs.center.x; e.p[1] = bs.center.y; e.p[2] = bs.center.z; e.p[3] = bs.radius;
			const uint32_t size = s->readUInt();                /* the OpenEXR file, decoded by the host's Bitmap */
			ref<MemoryStream> file = new MemoryStream(size);
			s->copyTo(file, size);
			file->setPos(0);
			env.bitmap = new Bitmap(Bitmap::EEXR, file);
			for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { env.l2w[4 * i + j] = l2w.m[i][j]; env.w2l[4 * i + j] = w2l.m[i][j]; }
		} else {
			SLog(EError, "gpupath: luminaire class \"%s\" is outside the supported hot path "
				"(area, point, spot, directional, constant, envmap, collimated)", cls.c_str());
		}
		(void) l2w;
	}

	/* Serialised through the manager, so that the BSDF itself has an id (1): textures that are XML children of the
	 * BSDF carry it as their parent and then write just that id instead of the whole BSDF again. */
	/* Analytic sphere: Shape::serialize (shape.cpp:130-138) + Sphere::serialize (sphere.cpp:71-77). The parent link (the
	 * scene) is cleared for the call as for the luminaires; the BSDF and the area luminaire that follow inline are read
	 * and dropped (the material comes from decodeBSDF, the luminaire from Scene::getLuminaires()). */
	struct SphereData { float o2w[16], w2o[16], radius, center[3]; int inverted; };
	static SphereData decodeSphere(const Shape *shapeConst) {
		Shape *shape = const_cast<Shape *>(shapeConst);
		ConfigurableObject *parent = shape->getParent();
		ref<MemoryStream> s = new MemoryStream();
		ref<InstanceManager> mgr = new InstanceManager();
		shape->setParent(NULL);
		mgr->serialize(s, shape);
		shape->setParent(parent);
		s->setPos(0);
		std::vector<mtsgpu_texture> scratch;
		std::vector<BitmapUpload> scratchBitmaps;
		Decoder d; d.s = s; d.textures = &scratch; d.bitmaps = &scratchBitmaps;
		d.maxId = s->readUInt();
		s->readString();
		readParent(d);
		{   /* m_bsdf */
			const unsigned int id = s->readUInt();
			if (id > d.maxId) { d.maxId = id; const std::string cls = s->readString(); decodePayload(cls, d); }
		}
		if (s->readUInt() != 0) SLog(EError, "gpupath: subsurface scattering is outside the supported hot path");
		{   /* m_luminaire: an AreaLuminaire written inline (luminaire.cpp:66-74, area.cpp:52-57); its values come through
		     * getLuminaires(), so it is only skipped here. Its parent and shape references are this sphere's id. */
			const unsigned int id = s->readUInt();
			if (id > d.maxId) {
				d.maxId = id;
				if (s->readString() != "AreaLuminaire") SLog(EError, "gpupath: unexpected luminaire class on a sphere");
				s->readUInt();                               /* parent: the sphere, already written */
				if (s->readUInt() != 0) SLog(EError, "gpupath: luminaires inside participating media are outside the supported hot path");
				s->readFloat(); s->readInt(); s->readBool();
				Transform worldToLuminaire(s);
				s->readString();
				Spectrum intensity(s);
				s->readUInt();                               /* m_shape: the sphere again */
			}
		}
		if (s->readUInt() != 0 || s->readUInt() != 0) SLog(EError, "gpupath: participating media are outside the supported hot path");
		s->readBool();                                   /* m_occluder */
		SphereData out;
		Transform objectToWorld(s);
		const Matrix4x4 &m = objectToWorld.getMatrix(), &mi = objectToWorld.getInverseMatrix();
		for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { out.o2w[4 * i + j] = m.m[i][j]; out.w2o[4 * i + j] = mi.m[i][j]; }
		out.radius = s->readFloat();
		Point c(s); out.center[0] = c.x; out.center[1] = c.y; out.center[2] = c.z;
		out.inverted = s->readBool() ? 1 : 0;
		return out;
	}

	/* Analytic cylinder: Shape::serialize + Cylinder::serialize (cylinder.cpp:69-74): objectToWorld, radius, length */
	static SphereData decodeCylinder(const Shape *shapeConst, float &length) {
		Shape *shape = const_cast<Shape *>(shapeConst);
		ConfigurableObject *parent = shape->getParent();
		ref<MemoryStream> s = new MemoryStream();
		ref<InstanceManager> mgr = new InstanceManager();
		shape->setParent(NULL);
		mgr->serialize(s, shape);
		shape->setParent(parent);
		s->setPos(0);
		std::vector<mtsgpu_texture> scratch;
		std::vector<BitmapUpload> scratchBitmaps;
		Decoder d; d.s = s; d.textures = &scratch; d.bitmaps = &scratchBitmaps;
		d.maxId = s->readUInt();
		s->readString();
		readParent(d);
		{   /* m_bsdf */
			const unsigned int id = s->readUInt();
			if (id > d.maxId) { d.maxId = id; const std::string cls = s->readString(); decodePayload(cls, d); }
		}
		if (s->readUInt() != 0) SLog(EError, "gpupath: subsurface scattering is outside the supported hot path");
		if (s->readUInt() != 0) SLog(EError, "gpupath: area luminaires on cylinders are outside the supported hot path");
		if (s->readUInt() != 0 || s->readUInt() != 0) SLog(EError, "gpupath: participating media are outside the supported hot path");
		s->readBool();                                   /* m_occluder */
		SphereData out; memset(&out, 0, sizeof(out));
		Transform objectToWorld(s);
		const Matrix4x4 &m = objectToWorld.getMatrix(), &mi = objectToWorld.getInverseMatrix();
		for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { out.o2w[4 * i + j] = m.m[i][j]; out.w2o[4 * i + j] = mi.m[i][j]; }
		out.radius = s->readFloat();
		length = s->readFloat();
		return out;
	}

	static mtsgpu_material decodeBSDF(const BSDF *bsdf, std::vector<mtsgpu_texture> &textures, std::vector<BitmapUpload> &bitmaps) {
		ref<MemoryStream> s = new MemoryStream();
		ref<InstanceManager> mgr = new InstanceManager();
		mgr->serialize(s, bsdf);
		s->setPos(0);
		Decoder d; d.s = s; d.maxId = 0; d.textures = &textures; d.bitmaps = &bitmaps;
		d.maxId = s->readUInt();
		const std::string cls = s->readString();
		return decodePayload(cls, d);
	}

	bool renderImpl(std::vector<mtsgpu_ctx *> &ctxs, Scene *scene, RenderQueue *queue, const RenderJob *job) {
		mtsgpu_ctx *ctx = ctxs[0];
		ref<Timer> timer = new Timer();
		const Camera *camera = scene->getCamera();
		Film *film = scene->getFilm();
		const Sampler *sampler = scene->getSampler();

		/* ---- refuse what the hot path does not cover ---- */
		if (!camera->getClass()->derivesFrom(MTS_CLASS(PerspectiveCamera)))
			Log(EError, "gpupath: only the perspective camera is supported");
		if (!scene->getMedia().empty() || camera->getMedium() != NULL)
			Log(EError, "gpupath: participating media are outside the supported hot path");

		/* ---- shapes in kd-tree order, materials, emitters ---- */
		const std::vector<const Shape *> &shapes = scene->getKDTree()->getShapes();
		const std::vector<Luminaire *> &luminaires = scene->getLuminaires();
		std::vector<mtsgpu_material> materials;
		std::vector<mtsgpu_texture> textures;
		std::vector<BitmapUpload> bitmaps;
		std::map<const BSDF *, int> materialIds;
		std::vector<mtsgpu_emitter> emitters(luminaires.size());
		std::map<const Luminaire *, int> emitterIds;
		EnvUpload env; int envIndex = -1;
		std::vector<EnvUpload> beams; std::vector<int> beamIndex;      /* collimated beams: their transforms */
		for (size_t i = 0; i < luminaires.size(); ++i) {
			const Luminaire *lum = luminaires[i];
			memset(&emitters[i], 0, sizeof(mtsgpu_emitter));
			emitters[i].samplingWeight = lum->getSamplingWeight();
			emitterIds[lum] = (int) i;
			if (lum->getClass()->getName() == "AreaLuminaire") {
				/* AreaLuminaire::Le returns m_intensity for any direction on the normal's side (area.cpp:62-66) */
				emitters[i].type = MTSGPU_EMITTER_AREA;
				put(emitters[i].intensity, lum->Le(ShapeSamplingRecord(Point(0.0f), Normal(0, 0, 1)), Vector(0, 0, 1)));
			} else {
				EnvUpload one;
				decodeLuminaire(lum, emitters[i], one);
				if (emitters[i].type == MTSGPU_EMITTER_ENVMAP) { envIndex = (int) i; env = one; }
				if (emitters[i].type == MTSGPU_EMITTER_COLLIMATED) { beamIndex.push_back((int) i); beams.push_back(one); }
			}
		}
		struct MeshRef { const TriMesh *mesh; int mat, emit; bool sphere; SphereData sph; int group; bool cylinder; float length; };   /* group >= 0: an instance (transform in sph) */
		struct GroupRef { std::vector<MeshRef> members; };
		std::vector<MeshRef> meshes;
		std::vector<GroupRef> groups;
		std::map<const ShapeGroup *, int> groupIds;
		for (size_t i = 0; i < shapes.size(); ++i) {
			const Shape *shape = shapes[i];
			const bool isSphere = shape->getClass()->getName() == "Sphere";
			const bool isCylinder = shape->getClass()->getName() == "Cylinder";
			if (shape->getClass()->getName() == "Instance") {
				/* instance.h:42-48: the group it shows (its kd-tree's meshes are uploaded once per group) and its transform */
				const Instance *inst = static_cast<const Instance *>(shape);
				const ShapeGroup *group = inst->getShapeGroup();
				if (!groupIds.count(group)) {
					GroupRef gr;
					const std::vector<const Shape *> &members = group->getKDTree()->getShapes();
					for (size_t k = 0; k < members.size(); ++k) {
						if (!members[k]->getClass()->derivesFrom(MTS_CLASS(TriMesh)))
							Log(EError, "gpupath: shape groups may hold triangle meshes only (found %s)", members[k]->getClass()->getName().c_str());
						MeshRef r; r.mesh = static_cast<const TriMesh *>(members[k]); r.mat = -1; r.emit = -1; r.sphere = false; r.group = -1; r.cylinder = false; r.length = 0;
						if (members[k]->hasBSDF()) {
							const BSDF *bsdf = members[k]->getBSDF();
							if (!materialIds.count(bsdf)) {
								materialIds[bsdf] = (int) materials.size();
								materials.push_back(decodeBSDF(bsdf, textures, bitmaps));
							}
							r.mat = materialIds[bsdf];
						}
						gr.members.push_back(r);
					}
					groupIds[group] = (int) groups.size();
					groups.push_back(gr);
				}
				MeshRef r; r.mesh = NULL; r.mat = -1; r.emit = -1; r.sphere = false; r.group = groupIds[group]; r.cylinder = false; r.length = 0;
				const Transform xf = inst->getWorldTransform();
				const Matrix4x4 &m = xf.getMatrix(), &mi = xf.getInverseMatrix();
				for (int a = 0; a < 4; ++a) for (int b = 0; b < 4; ++b) { r.sph.o2w[4 * a + b] = m.m[a][b]; r.sph.w2o[4 * a + b] = mi.m[a][b]; }
				meshes.push_back(r);
				continue;
			}
			if (!isSphere && !isCylinder && !shape->getClass()->derivesFrom(MTS_CLASS(TriMesh)))
				Log(EError, "gpupath: shape %i (%s) is neither a triangle mesh, a sphere, a cylinder nor an instance", (int) i, shape->getClass()->getName().c_str());
			if (shape->hasSubsurface() || shape->isMediumTransition())
				Log(EError, "gpupath: subsurface / medium transitions are outside the supported hot path");
			MeshRef r; r.mesh = (isSphere || isCylinder) ? NULL : static_cast<const TriMesh *>(shape); r.mat = -1; r.emit = -1; r.sphere = isSphere; r.group = -1;
			r.cylinder = isCylinder; r.length = 0;
			if (isSphere) r.sph = decodeSphere(shape);
			if (isCylinder) r.sph = decodeCylinder(shape, r.length);
			if (shape->hasBSDF()) {
				const BSDF *bsdf = shape->getBSDF();
				if (!materialIds.count(bsdf)) {
					materialIds[bsdf] = (int) materials.size();
					materials.push_back(decodeBSDF(bsdf, textures, bitmaps));
				}
				r.mat = materialIds[bsdf];
			}
			if (shape->isLuminaire())
				r.emit = emitterIds[shape->getLuminaire()];
			meshes.push_back(r);
		}
		for (size_t g = 0; g < ctxs.size(); ++g) {       /* the scene is replicated on every GPU */
			mtsgpu_ctx *c = ctxs[g];
			check(c, mtsgpu_upload_textures(c, textures.empty() ? NULL : &textures[0], (int) textures.size()));
			for (size_t i = 0; i < bitmaps.size(); ++i) {
				const Bitmap *bmp = bitmaps[i].bitmap;
				const int bpp = bmp->getBitsPerPixel(), w = bmp->getWidth(), h = bmp->getHeight();
				if (bpp == 128) {                   /* an EXRTexture's floats */
					check(c, mtsgpu_upload_bitmap(c, bitmaps[i].texture, w, h, 4, bmp->getData(), MTSGPU_BITMAP_F32, 1.0f));
				} else if (bpp == 1) {              /* ldrtexture.cpp:185-199: bit k of byte k / 8, set -> 255 */
					std::vector<uint8_t> bytes((size_t) w * h);
					for (size_t k = 0; k < bytes.size(); ++k) bytes[k] = (bmp->getData()[k / 8] & (1 << (k % 8))) ? 255 : 0;
					check(c, mtsgpu_upload_bitmap(c, bitmaps[i].texture, w, h, 1, &bytes[0], MTSGPU_BITMAP_U8, bitmaps[i].gamma));
				} else {
					check(c, mtsgpu_upload_bitmap(c, bitmaps[i].texture, w, h, bpp / 8, bmp->getData(), MTSGPU_BITMAP_U8, bitmaps[i].gamma));
				}
			}
			check(c, mtsgpu_upload_materials(c, materials.empty() ? NULL : &materials[0], (int) materials.size()));
			check(c, mtsgpu_upload_emitters(c, emitters.empty() ? NULL : &emitters[0], (int) emitters.size()));
			for (size_t b = 0; b < beams.size(); ++b)
				check(c, mtsgpu_upload_emitter_transform(c, beamIndex[b], beams[b].l2w, beams[b].w2l));
			if (envIndex >= 0)      /* Bitmap::loadEXR leaves RGBA floats (bitmap.cpp:247-273) */
				check(c, mtsgpu_upload_envmap(c, envIndex, env.bitmap->getWidth(), env.bitmap->getHeight(), 4,
					env.bitmap->getFloatData(), env.l2w, env.w2l));
			for (size_t gi = 0; gi < groups.size(); ++gi)
				for (size_t k = 0; k < groups[gi].members.size(); ++k) {
					const TriMesh *m = groups[gi].members[k].mesh;
					check(c, mtsgpu_upload_group_mesh(c, (int) gi, (int) k,
						reinterpret_cast<const float *>(m->getVertexPositions()),
						reinterpret_cast<const float *>(m->getVertexNormals()),
						reinterpret_cast<const float *>(m->getVertexTexcoords()), (uint32_t) m->getVertexCount(),
						reinterpret_cast<const uint32_t *>(m->getTriangles()), (uint32_t) m->getTriangleCount(), groups[gi].members[k].mat));
				}
			for (size_t i = 0; i < meshes.size(); ++i) {
				if (meshes[i].group >= 0) {
					check(c, mtsgpu_upload_instance(c, (int) i, meshes[i].group, meshes[i].sph.o2w, meshes[i].sph.w2o));
					continue;
				}
				if (meshes[i].cylinder) {
					check(c, mtsgpu_upload_cylinder(c, (int) i, meshes[i].sph.o2w, meshes[i].sph.w2o, meshes[i].sph.radius, meshes[i].length, meshes[i].mat));
					continue;
				}
				if (meshes[i].sphere) {
					const SphereData &sp = meshes[i].sph;
					check(c, mtsgpu_upload_sphere(c, (int) i, sp.o2w, sp.w2o, sp.radius, sp.center, sp.inverted, meshes[i].mat, meshes[i].emit));
					continue;
				}
				const TriMesh *m = meshes[i].mesh;
				/* Point/Normal are 3 packed floats, Triangle is 3 packed uint32 (core/triangle.h) */
				check(c, mtsgpu_upload_mesh(c, (int) i,
					reinterpret_cast<const float *>(m->getVertexPositions()),
					reinterpret_cast<const float *>(m->getVertexNormals()),
					reinterpret_cast<const float *>(m->getVertexTexcoords()), (uint32_t) m->getVertexCount(),
					reinterpret_cast<const uint32_t *>(m->getTriangles()), (uint32_t) m->getTriangleCount(),
					meshes[i].mat, meshes[i].emit));
			}
		}

		/* ---- camera: m_rasterToCamera is private to the plugin class, so rebuild it with the same
		        Transform operations as PerspectiveCameraImpl::configure (perspective.cpp:43-71) ---- */
		const PerspectiveCamera *pc = static_cast<const PerspectiveCamera *>(camera);
		const Vector2i fsize = film->getSize();
		const Float aspect = (Float) fsize.x / (Float) fsize.y;
		const bool mapSmallerSide = camera->getProperties().getBoolean("mapSmallerSide", true);
		Transform screenToRaster;
		if ((aspect >= 1.0f) ^ !mapSmallerSide) {
			screenToRaster = Transform::scale(Vector((Float) fsize.x, (Float) fsize.y, 1.0f))
				* Transform::scale(Vector(1 / (2 * aspect), -0.5f, 1.0f))
				* Transform::translate(Vector(aspect, -1.0f, 0));
		} else {
			screenToRaster = Transform::scale(Vector((Float) fsize.x, (Float) fsize.y, 1.0f))
				* Transform::scale(Vector(0.5f, -0.5f * aspect, 1.0f))
				* Transform::translate(Vector(1.0f, -1 / aspect, 0));
		}
		const Transform rasterToCamera = pc->getProjectionTransform().inverse() * screenToRaster.inverse();
		{
			/* cross-check against the camera's own generateRay at three raster positions */
			const Point2 probes[3] = { Point2(0.5f, 0.5f), Point2(fsize.x - 0.5f, 0.5f), Point2(fsize.x * 0.5f, fsize.y - 0.5f) };
			for (int k = 0; k < 3 && pc->getApertureRadius() == 0; ++k) {
				Ray r; camera->generateRay(probes[k], Point2(0.5f), 0.0f, r);
				Point ic = rasterToCamera(Point(probes[k].x, probes[k].y, 0));
				Vector d = camera->getInverseViewTransform()(normalize(Vector(ic)));
				if ((d - r.d).length() > 1e-4f)
					Log(EError, "gpupath: could not reproduce the camera's raster-to-camera transform");
			}
		}
		float r2c[16], c2w[16];
		for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) {
			r2c[i * 4 + j] = rasterToCamera.getMatrix().m[i][j];
			c2w[i * 4 + j] = camera->getInverseViewTransform().getMatrix().m[i][j];
		}
		for (size_t g = 0; g < ctxs.size(); ++g)
			check(ctxs[g], mtsgpu_set_camera(ctxs[g], r2c, c2w, pc->getNearClip(), pc->getFarClip(), pc->getApertureRadius(),
				pc->getFocusDepth(), camera->getShutterOpen(), camera->getShutterOpenTime()));

		/* ---- film + tabulated filter (rfilter.h:65-118) ---- */
		const TabulatedFilter *filter = film->getTabulatedFilter();
		float table[256];
		for (int y = 0; y < 16; ++y) for (int x = 0; x < 16; ++x) table[y * 16 + x] = filter->lookup(x, y);
		const Vector2i csize = film->getCropSize(); const Point2i coff = film->getCropOffset();
		for (size_t g = 0; g < ctxs.size(); ++g) {
			check(ctxs[g], mtsgpu_set_film(ctxs[g], fsize.x, fsize.y, coff.x, coff.y, csize.x, csize.y, table,
				filter->getFilterSize().x, filter->getFilterSize().y));
			check(ctxs[g], mtsgpu_build_accel(ctxs[g]));
		}
		mtsgpu_accel_info info; check(ctx, mtsgpu_accel_info_get(ctx, &info));
		Log(EInfo, "gpupath: %u triangles, %u BVH nodes, built on the GPU in %.1f ms (upload + build %i ms)",
			(unsigned) info.triangles, (unsigned) info.nodes, info.buildMs, timer->getMilliseconds());

		/* ---- render ---- */
		mtsgpu_render_params P; memset(&P, 0, sizeof(P));
		P.mode = m_mode; P.maxDepth = m_maxDepth; P.rrDepth = m_rrDepth; P.strictNormals = m_strictNormals ? 1 : 0;
		P.luminaireSamples = m_luminaireSamples; P.bsdfSamples = m_bsdfSamples;
		P.sppTotal = (uint32_t) sampler->getSampleCount(); P.sppBegin = 0; P.sppEnd = P.sppTotal;
		P.seed = (uint32_t) m_seed;
		/* Scene property importanceSampleLuminaires (scene.cpp:51): it selects the luminaire CDF (scene.cpp:320-330) and the
		   fraction in Scene::pdfLuminaire (scene.cpp:381-394). The member is private and has no getter; it only changes
		   anything when the luminaires' sampling weights differ, and then Scene::toString() states it (scene.cpp:759). */
		P.importanceSampleLuminaires = 1;
		{
			const std::vector<Luminaire *> &lums = scene->getLuminaires();
			bool weightsDiffer = false;
			for (size_t i = 1; i < lums.size(); ++i)
				weightsDiffer |= lums[i]->getSamplingWeight() != lums[0]->getSamplingWeight();
			if (weightsDiffer) {
				const std::string desc = scene->toString();
				const std::string key = "importanceSampleLuminaires = ";
				const size_t pos = desc.find(key);
				if (pos == std::string::npos || pos + key.size() >= desc.size())
					Log(EError, "gpupath: cannot tell whether the scene importance-samples its luminaires");
				P.importanceSampleLuminaires = desc[pos + key.size()] != '0' ? 1 : 0;
			}
		}
		P.samplerType = MTSGPU_SAMPLER_INDEPENDENT; P.samplerDepth = 0;
		{
			const std::string samplerClass = sampler->getClass()->getName();
			if (samplerClass == "LowDiscrepancySampler" || samplerClass == "StratifiedSampler") {
				/* m_depth has no getter: read it from LowDiscrepancySampler::serialize (ldsampler.cpp:76-80), which writes
				   Sampler::serialize (sampler.cpp:39-49) first; the parent link (the camera) is cleared for the call */
				Sampler *smp = const_cast<Sampler *>(sampler);
				ConfigurableObject *parent = smp->getParent();
				ref<MemoryStream> ms = new MemoryStream();
				ref<InstanceManager> mgr = new InstanceManager();
				smp->setParent(NULL);
				smp->serialize(ms, mgr);
				smp->setParent(parent);
				ms->setPos(0);
				if (ms->readUInt() != 0) Log(EError, "gpupath: unexpected sampler parent");
				ms->readSize();
				for (size_t k = ms->readSize(); k > 0; --k) ms->readUInt();
				for (size_t k = ms->readSize(); k > 0; --k) ms->readUInt();
				/* both classes write Sampler::serialize, then m_depth (ldsampler.cpp:76-80, stratified.cpp:84-90) */
				P.samplerType = samplerClass == "StratifiedSampler" ? MTSGPU_SAMPLER_STRATIFIED : MTSGPU_SAMPLER_LD;
				P.samplerDepth = (uint32_t) ms->readInt();
			} else if (samplerClass == "HaltonSequence") {
				P.samplerType = MTSGPU_SAMPLER_HALTON;
			} else if (samplerClass == "HammersleySequence") {
				P.samplerType = MTSGPU_SAMPLER_HAMMERSLEY;
			} else if (samplerClass != "IndependentSampler") {
				Log(EError, "gpupath: sampler class \"%s\" is outside the supported hot path (independent, ldsampler, stratified, halton, hammersley)", samplerClass.c_str());
			}
		}
		const size_t nPix = (size_t) csize.x * csize.y;
		std::vector<float> accum(nPix * 5);
		mtsgpu_stats stats;
		ProgressReporter progress("Rendering", (long long) nPix * (long long) P.sppTotal, job);
		RefreshState refresh = { &progress, film, queue, job, csize, coff, (long long) nPix * (long long) P.sppTotal, 0 };
		/* previews need the whole film on one GPU; with several GPUs only the progress bar moves */
		check(ctx, mtsgpu_set_progress(ctx, &GPUPathIntegrator::onProgress, &refresh, m_refreshInterval, ctxs.size() == 1 ? 1 : 0));
		timer->reset();
		const int rc = ctxs.size() == 1 ? mtsgpu_render(ctx, &P, &m_cancel, &accum[0], &stats)
			: mtsgpu_render_multi(&ctxs[0], (int) ctxs.size(), &P, &m_cancel, &accum[0], &stats);   /* sample range split + NVLink peer reduce */
		mtsgpu_set_progress(ctx, NULL, NULL, 0.0, 0);
		if (rc == MTSGPU_ERR_CANCELLED)
			return false;
		check(ctx, rc);
		gpuNormalRays += (size_t) stats.normalRays; gpuShadowRays += (size_t) stats.shadowRays;
		Log(EInfo, "gpupath: %.1f M samples in %i ms (%.1f Msamples/s), %.1f M closest-hit + %.1f M shadow rays "
			"(closest %.1f ms, shadow %.1f ms on the device)", stats.samples * 1e-6, timer->getMilliseconds(),
			stats.samples * 1e-3 / std::max(stats.totalMs, 1e-3), stats.normalRays * 1e-6, stats.shadowRays * 1e-6,
			stats.traceClosestMs, stats.traceAnyMs);

		if (refresh.refreshes > 0)
			film->clear();            /* the previews were put into the film; the final accumulators replace them */
		ref<ImageBlock> block = makeBlock(&accum[0], csize, coff);
		film->putImageBlock(block);
		progress.finish();
		queue->signalWorkEnd(job, block);
		return true;
	}

	/* {sum spec, sum alpha, sum weight} as one full-frame ImageBlock, through the block's own wire format
	   (ImageBlock::load, imageblock.cpp:89-106): there is no alpha setter */
	static ref<ImageBlock> makeBlock(const float *accum, const Vector2i &csize, const Point2i &coff) {
		const size_t nPix = (size_t) csize.x * csize.y;
		ref<ImageBlock> block = new ImageBlock(csize, 0, true, true, false, false);
		ref<MemoryStream> ms = new MemoryStream(nPix * 20 + 64);
		coff.serialize(ms); csize.serialize(ms);
		std::vector<Float> tmp(nPix * 3);
		for (size_t i = 0; i < nPix; ++i) {
			Spectrum s; s.fromLinearRGB(accum[i * 5], accum[i * 5 + 1], accum[i * 5 + 2]);
			Float r, g, b; s.toLinearRGB(r, g, b);
			tmp[i * 3] = r; tmp[i * 3 + 1] = g; tmp[i * 3 + 2] = b;
		}
		ms->writeFloatArray(&tmp[0], nPix * 3);
		for (size_t i = 0; i < nPix; ++i) tmp[i] = accum[i * 5 + 3];
		ms->writeFloatArray(&tmp[0], nPix);
		for (size_t i = 0; i < nPix; ++i) tmp[i] = accum[i * 5 + 4];
		ms->writeFloatArray(&tmp[0], nPix);
		ms->writeInt(0);
		ms->setPos(0);
		block->load(ms);
		return block;
	}

	/* Progress bar and preview while the GPU renders: what BlockedRenderProcess::processResult does after every block
	   (renderproc.cpp:123-130: merge into the film, ProgressReporter::update, and the GUI / `-r` path through
	   RenderQueue::signalRefresh, renderqueue.h:109-112). Called by the library between wavefront batches, on this thread. */
	struct RefreshState {
		ProgressReporter *progress; Film *film; RenderQueue *queue; const RenderJob *job;
		Vector2i csize; Point2i coff; long long total; int refreshes;
	};
	static void onProgress(void *user, double fraction, const float *accum) {
		RefreshState *r = static_cast<RefreshState *>(user);
		r->progress->update((long long) (fraction * (double) r->total));
		if (accum) {
			r->film->clear();
			ref<ImageBlock> block = makeBlock(accum, r->csize, r->coff);
			r->film->putImageBlock(block);
			r->queue->signalRefresh(r->job, NULL);
			r->refreshes++;
		}
	}

	int m_mode, m_maxDepth, m_rrDepth, m_luminaireSamples, m_bsdfSamples, m_device, m_seed, m_gpus;
	bool m_strictNormals;
	Float m_refreshInterval;      /* seconds between previews (property `refreshInterval`, default 2) */
	volatile int m_cancel;
};

MTS_IMPLEMENT_CLASS_S(GPUPathIntegrator, false, SampleIntegrator)
MTS_EXPORT_PLUGIN(GPUPathIntegrator, "CUDA (sm_100a) wavefront path tracer");
MTS_NAMESPACE_END

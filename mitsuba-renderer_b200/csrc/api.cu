// C ABI entry points of libmtscuda.so (declared in include/mtsgpu.h).
#include "context.h"
#include <omp.h>
#include <cstring>
#include <cmath>
#include <algorithm>

int setError(mtsgpu_ctx *ctx, int code, const char *fmt, ...) {
	char buf[1024];
	va_list ap; va_start(ap, fmt);
	vsnprintf(buf, sizeof(buf), fmt, ap);
	va_end(ap);
	if (ctx) ctx->error = buf;
	return code;
}

static thread_local std::string g_createError;

void setGlobalError(const char *fmt, ...) {      // context-free calls (mtsgpu_create, mtsgpu_serialized_*)
	char buf[1024];
	va_list ap; va_start(ap, fmt);
	vsnprintf(buf, sizeof(buf), fmt, ap);
	va_end(ap);
	g_createError = buf;
}

// The library's own copy of a caller's array ("the library copies during the call"), on all host cores for large ones:
// a single-threaded vector::assign of the 216 MB of a 10M-triangle scene took 96 ms, mostly page faults.
template <typename V, typename T> static void hostCopy(V &dst, const T *src, size_t n) {
	dst.resize(n);                                   // NoInitAlloc: no value-initialisation pass
	const size_t chunk = (size_t) 1 << 18;           // elements
	const int64_t nChunks = (int64_t) ((n + chunk - 1) / chunk);
#pragma omp parallel for schedule(static) num_threads(hostThreads()) if (nChunks > 4)
	for (int64_t c = 0; c < nChunks; ++c) {
		const size_t b = (size_t) c * chunk, e = std::min(n, b + chunk);
		memcpy(dst.data() + b, src + b, (e - b) * sizeof(T));
	}
}

// Threads for the host-side copies and flatten loops: MTSGPU_HOST_THREADS, else all cores (at most 16). Not OMP_NUM_THREADS:
// launchers that start one process per GPU set it to 1 for every rank (torchrun does), which made upload + build 2.4x slower.
int hostThreads() {
	static const int n = [] {
		const char *e = getenv("MTSGPU_HOST_THREADS");
		const int want = e ? atoi(e) : std::min(omp_get_num_procs(), 16);
		return std::min(std::max(want, 1), 64);
	}();
	return n;
}

extern "C" {

const char *mtsgpu_version(void) { return "mtsgpu 0.1 (sm_100a)"; }

int mtsgpu_create(int device, mtsgpu_ctx **out) {
	if (!out) return MTSGPU_ERR_ARG;
	*out = nullptr;
	int count = 0;
	cudaError_t e = cudaGetDeviceCount(&count);
	if (e != cudaSuccess || count == 0) {
		g_createError = std::string("no usable CUDA device: ") + cudaGetErrorString(e);
		return MTSGPU_ERR_CUDA;      // no CPU fallback by design
	}
	if (device < 0 || device >= count) { g_createError = "device index out of range"; return MTSGPU_ERR_ARG; }
	if ((e = cudaSetDevice(device)) != cudaSuccess) { g_createError = cudaGetErrorString(e); return MTSGPU_ERR_CUDA; }
	mtsgpu_ctx *ctx = new mtsgpu_ctx();
	ctx->device = device;
	memset(&ctx->scene, 0, sizeof(ctx->scene));
	if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) {
		g_createError = cudaGetErrorString(e); delete ctx; return MTSGPU_ERR_CUDA;
	}
	for (int i = 0; i < 8; ++i)
		if ((e = cudaEventCreate(&ctx->ev[i])) != cudaSuccess) {
			g_createError = std::string("cudaEventCreate: ") + cudaGetErrorString(e);
			for (int k = 0; k < i; ++k) cudaEventDestroy(ctx->ev[k]);
			cudaStreamDestroy(ctx->stream); delete ctx; return MTSGPU_ERR_CUDA;
		}
	*out = ctx;
	return MTSGPU_OK;
}

void mtsgpu_destroy(mtsgpu_ctx *ctx) {
	if (!ctx) return;
	for (mtsgpu_ctx *g : ctx->groups) mtsgpu_destroy(g);
	ctx->groups.clear();
	cudaSetDevice(ctx->device);
	cudaStreamSynchronize(ctx->stream);
	for (int i = 0; i < 8; ++i) if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
	if (ctx->copyIn) {
		cudaStreamDestroy(ctx->copyIn); cudaStreamDestroy(ctx->copyOut);
		for (int i = 0; i < 2 * MTSGPU_PIPE_CHUNKS; ++i) cudaEventDestroy(ctx->pipeEv[i]);
	}
	cudaStreamDestroy(ctx->stream);
	if (ctx->hCounts) cudaFreeHost(ctx->hCounts);
	if (ctx->hProgressFilm) cudaFreeHost(ctx->hProgressFilm);
	if (ctx->progressStream) { cudaStreamDestroy(ctx->progressStream); cudaEventDestroy(ctx->progressSnap); cudaEventDestroy(ctx->progressCopied); }
	for (int i = 0; i < 16; ++i) if (ctx->pollEv[i]) cudaEventDestroy(ctx->pollEv[i]);
	for (int k = 0; k < 2; ++k) for (cudaEvent_t e : ctx->evPool[k]) cudaEventDestroy(e);
	delete ctx;
}

const char *mtsgpu_last_error(const mtsgpu_ctx *ctx) {
	return ctx ? ctx->error.c_str() : g_createError.c_str();
}

#define ENTER(ctx) do { if (!(ctx)) return MTSGPU_ERR_ARG; cudaError_t _e = cudaSetDevice((ctx)->device); \
	if (_e != cudaSuccess) return setError((ctx), MTSGPU_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(_e)); } while (0)

int mtsgpu_upload_materials(mtsgpu_ctx *ctx, const mtsgpu_material *materials, int count) {
	ENTER(ctx);
	REQUIRE(ctx, count >= 0 && (count == 0 || materials), MTSGPU_ERR_ARG, "upload_materials: bad arguments");
	for (int i = 0; i < count; ++i) {
		const uint32_t t = materials[i].type;
		REQUIRE(ctx, (t & MTSGPU_BSDF_TYPE_MASK) <= MTSGPU_BSDF_TYPE_LAST && !(t & ~(MTSGPU_BSDF_TYPE_MASK | MTSGPU_BSDF_TWOSIDED | MTSGPU_BSDF_MASKED)),
			MTSGPU_ERR_UNSUPPORTED, "material %d: unsupported BSDF type 0x%x", i, t);
		REQUIRE(ctx, !((t & MTSGPU_BSDF_TWOSIDED) && ((t & MTSGPU_BSDF_TYPE_MASK) == MTSGPU_BSDF_DIELECTRIC ||
			(t & MTSGPU_BSDF_TYPE_MASK) == MTSGPU_BSDF_ROUGHGLASS)), MTSGPU_ERR_UNSUPPORTED,
			"material %d: TwoSidedBRDF: only BRDF child instances (without transmission) are supported", i);
	}
	ctx->materials.assign(materials, materials + count);
	ctx->built = false;
	for (mtsgpu_ctx *g : ctx->groups) g->built = false;      // groups share these tables: rebuild (and re-validate) them too
	return MTSGPU_OK;
}

int mtsgpu_upload_textures(mtsgpu_ctx *ctx, const mtsgpu_texture *textures, int count) {
	ENTER(ctx);
	REQUIRE(ctx, count >= 0 && (count == 0 || textures), MTSGPU_ERR_ARG, "upload_textures: bad arguments");
	for (int i = 0; i < count; ++i)
		REQUIRE(ctx, textures[i].type == MTSGPU_TEXTURE_CHECKERBOARD || textures[i].type == MTSGPU_TEXTURE_BITMAP || textures[i].type == MTSGPU_TEXTURE_GRID, MTSGPU_ERR_UNSUPPORTED,
			"texture %d: unsupported texture type %u", i, textures[i].type);
	for (int i = 0; i < count; ++i) {
		if (textures[i].type != MTSGPU_TEXTURE_BITMAP) continue;
		const float f = textures[i].p[0], w = textures[i].p[1];
		REQUIRE(ctx, (f == 0 || f == 1 || f == 2) && (w == 0 || w == 1 || w == 2 || w == 3), MTSGPU_ERR_ARG,
			"texture %d: unknown filter type / wrap mode (%g, %g)", i, f, w);
	}
	ctx->textures.assign(textures, textures + count);
	ctx->bitmaps.assign(count, HostBitmap());
	ctx->built = false;
	for (mtsgpu_ctx *g : ctx->groups) g->built = false;      // groups share these tables: rebuild (and re-validate) them too
	return MTSGPU_OK;
}

int mtsgpu_upload_bitmap(mtsgpu_ctx *ctx, int textureIndex, int width, int height, int channels, const void *pixels, int format, float gamma) {
	ENTER(ctx);
	REQUIRE(ctx, textureIndex >= 0 && (size_t) textureIndex < ctx->textures.size() && ctx->textures[textureIndex].type == MTSGPU_TEXTURE_BITMAP,
		MTSGPU_ERR_ARG, "upload_bitmap: texture %d is not a bitmap texture of the uploaded table", textureIndex);
	REQUIRE(ctx, pixels && width > 0 && height > 0 && width <= 16384 && height <= 16384, MTSGPU_ERR_ARG, "upload_bitmap: bad size %d x %d", width, height);
	HostBitmap &b = ctx->bitmaps[textureIndex];
	const size_t n = (size_t) width * height;
	if (format == MTSGPU_BITMAP_U8) {
		// "%i bpp images are currently not supported!" (ldrtexture.cpp:201)
		REQUIRE(ctx, channels == 1 || channels == 2 || channels == 3 || channels == 4, MTSGPU_ERR_ARG, "upload_bitmap: %d bpp images are currently not supported!", 8 * channels);
		float tbl[256];                                 // LDRTexture::initializeFrom (ldrtexture.cpp:117-133)
		for (int i = 0; i < 256; ++i) {
			const float value = (float) i / (float) 255;
			if (gamma == -1) tbl[i] = value <= 0.04045f ? value / 12.92f : std::pow((value + 0.055f) / (float) (1.0 + 0.055), 2.4f);
			else tbl[i] = std::pow(value, gamma);
		}
		const uint8_t *d = static_cast<const uint8_t *>(pixels);
		b.rgb.resize(3 * n);
		for (size_t i = 0; i < n; ++i)
			for (int c = 0; c < 3; ++c) b.rgb[3 * i + c] = tbl[d[i * channels + (channels >= 3 ? c : 0)]];
	} else {
		REQUIRE(ctx, format == MTSGPU_BITMAP_F32 && (channels == 3 || channels == 4), MTSGPU_ERR_ARG, "upload_bitmap: bad format / channel count");
		const float *d = static_cast<const float *>(pixels);
		b.rgb.resize(3 * n);
		for (size_t i = 0; i < n; ++i) for (int c = 0; c < 3; ++c) b.rgb[3 * i + c] = d[i * channels + c];
	}
	b.width = width; b.height = height;
	ctx->built = false;
	for (mtsgpu_ctx *g : ctx->groups) g->built = false;
	return MTSGPU_OK;
}

int mtsgpu_upload_emitters(mtsgpu_ctx *ctx, const mtsgpu_emitter *emitters, int count) {
	ENTER(ctx);
	REQUIRE(ctx, count >= 0 && (count == 0 || emitters), MTSGPU_ERR_ARG, "upload_emitters: bad arguments");
	int backgrounds = 0;
	for (int i = 0; i < count; ++i) {
		REQUIRE(ctx, emitters[i].type <= MTSGPU_EMITTER_COLLIMATED, MTSGPU_ERR_UNSUPPORTED, "emitter %d: unsupported luminaire type %u", i, emitters[i].type);
		if (emitters[i].type == MTSGPU_EMITTER_SPOT)      // Assert(m_cutoffAngle >= m_beamWidth), spot.cpp:42
			REQUIRE(ctx, emitters[i].p[7] >= emitters[i].p[6], MTSGPU_ERR_ARG, "emitter %d: cutoffAngle must be >= beamWidth", i);
		backgrounds += emitters[i].type == MTSGPU_EMITTER_CONSTANT || emitters[i].type == MTSGPU_EMITTER_ENVMAP;
	}
	REQUIRE(ctx, backgrounds <= 1, MTSGPU_ERR_UNSUPPORTED, "upload_emitters: at most one background luminaire");
	ctx->emitters.assign(emitters, emitters + count);
	ctx->envEmitter = -1; ctx->envBitmap = HostBitmap();
	ctx->emitterXf.assign((size_t) count * 32, 0.0f);
	ctx->built = false;
	return MTSGPU_OK;
}

int mtsgpu_upload_emitter_transform(mtsgpu_ctx *ctx, int emitterIndex, const float luminaireToWorld[16], const float worldToLuminaire[16]) {
	ENTER(ctx);
	REQUIRE(ctx, luminaireToWorld && worldToLuminaire && emitterIndex >= 0 && (size_t) emitterIndex < ctx->emitters.size(), MTSGPU_ERR_ARG,
		"upload_emitter_transform: emitter %d is not in the uploaded table", emitterIndex);
	memcpy(&ctx->emitterXf[(size_t) emitterIndex * 32], luminaireToWorld, 64);
	memcpy(&ctx->emitterXf[(size_t) emitterIndex * 32 + 16], worldToLuminaire, 64);
	ctx->lumImportance = -1;
	return MTSGPU_OK;
}

int mtsgpu_upload_envmap(mtsgpu_ctx *ctx, int emitterIndex, int width, int height, int channels, const float *pixels,
		const float luminaireToWorld[16], const float worldToLuminaire[16]) {
	ENTER(ctx);
	REQUIRE(ctx, emitterIndex >= 0 && (size_t) emitterIndex < ctx->emitters.size() && ctx->emitters[emitterIndex].type == MTSGPU_EMITTER_ENVMAP,
		MTSGPU_ERR_ARG, "upload_envmap: emitter %d is not an environment map of the uploaded table", emitterIndex);
	REQUIRE(ctx, pixels && luminaireToWorld && worldToLuminaire && width > 0 && height > 0 && width <= 16384 && height <= 16384 && (channels == 3 || channels == 4),
		MTSGPU_ERR_ARG, "upload_envmap: bad arguments");
	HostBitmap &b = ctx->envBitmap;
	const size_t n = (size_t) width * height;
	b.rgb.resize(3 * n);
	for (size_t i = 0; i < n; ++i) for (int c = 0; c < 3; ++c) b.rgb[3 * i + c] = pixels[i * channels + c];
	b.width = width; b.height = height;
	for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) { ctx->envL2W[3 * r + c] = luminaireToWorld[4 * r + c]; ctx->envW2L[3 * r + c] = worldToLuminaire[4 * r + c]; }
	ctx->envEmitter = emitterIndex;
	ctx->built = false;
	return MTSGPU_OK;
}

int mtsgpu_upload_mesh(mtsgpu_ctx *ctx, int shapeIndex, const float *positions, const float *normals, const float *texcoords,
		uint32_t nVert, const uint32_t *triangles, uint32_t nTri, int materialId, int emitterId) {
	ENTER(ctx);
	REQUIRE(ctx, positions && triangles && nVert > 0 && nTri > 0, MTSGPU_ERR_ARG, "upload_mesh: empty mesh (TriMesh::configure rejects these too)");
	REQUIRE(ctx, shapeIndex >= 0 && (size_t) shapeIndex <= ctx->meshes.size(), MTSGPU_ERR_ARG,
		"upload_mesh: shapes must be uploaded in kd-tree order (got %d, expected <= %zu)", shapeIndex, ctx->meshes.size());
	if ((size_t) shapeIndex < ctx->meshes.size()) ctx->meshes.resize(shapeIndex);   // re-upload from this shape on
	HostMesh m;
	m.nVert = nVert; m.nTri = nTri; m.materialId = materialId; m.emitterId = emitterId; m.hasNormals = normals != nullptr;
	hostCopy(m.pos, positions, 3 * (size_t) nVert);
	if (normals) hostCopy(m.nrm, normals, 3 * (size_t) nVert);
	if (texcoords) hostCopy(m.uv, texcoords, 2 * (size_t) nVert);
	hostCopy(m.tri, triangles, 3 * (size_t) nTri);
	ctx->meshes.push_back(std::move(m));
	ctx->built = false;
	return MTSGPU_OK;
}

int mtsgpu_upload_sphere(mtsgpu_ctx *ctx, int shapeIndex, const float objectToWorld[16], const float worldToObject[16],
		float radius, const float center[3], int inverted, int materialId, int emitterId) {
	ENTER(ctx);
	REQUIRE(ctx, objectToWorld && worldToObject && center, MTSGPU_ERR_ARG, "upload_sphere: null argument");
	REQUIRE(ctx, shapeIndex >= 0 && (size_t) shapeIndex <= ctx->meshes.size(), MTSGPU_ERR_ARG,
		"upload_sphere: shapes must be uploaded in kd-tree order (got %d, expected <= %zu)", shapeIndex, ctx->meshes.size());
	if ((size_t) shapeIndex < ctx->meshes.size()) ctx->meshes.resize(shapeIndex);
	HostMesh m;
	m.nVert = 1; m.nTri = 1; m.materialId = materialId; m.emitterId = emitterId < 0 ? -1 : emitterId; m.hasNormals = false;
	m.pos.assign(center, center + 3);
	m.tri = { 0u, SPHERE_MARK, SPHERE_MARK };
	m.sphere = true; m.inverted = inverted != 0; m.radius = radius;
	memcpy(m.o2w, objectToWorld, 64); memcpy(m.w2o, worldToObject, 64);
	ctx->meshes.push_back(std::move(m));
	ctx->built = false;
	return MTSGPU_OK;
}

int mtsgpu_upload_cylinder(mtsgpu_ctx *ctx, int shapeIndex, const float objectToWorld[16], const float worldToObject[16],
		float radius, float length, int materialId) {
	ENTER(ctx);
	REQUIRE(ctx, objectToWorld && worldToObject, MTSGPU_ERR_ARG, "upload_cylinder: null argument");
	REQUIRE(ctx, shapeIndex >= 0 && (size_t) shapeIndex <= ctx->meshes.size(), MTSGPU_ERR_ARG,
		"upload_cylinder: shapes must be uploaded in kd-tree order (got %d, expected <= %zu)", shapeIndex, ctx->meshes.size());
	REQUIRE(ctx, length > 0 && radius > 0, MTSGPU_ERR_ARG, "upload_cylinder: m_length > 0 && m_radius > 0 (cylinder.cpp:62)");
	if ((size_t) shapeIndex < ctx->meshes.size()) ctx->meshes.resize(shapeIndex);
	HostMesh m;
	m.nVert = 2; m.nTri = 1; m.materialId = materialId; m.emitterId = -1; m.hasNormals = false;
	m.pos.assign(6, 0.0f);
	m.tri = { 0u, INSTANCE_MARK, 1u };
	m.cylinder = true; m.radius = radius; m.length = length;
	memcpy(m.o2w, objectToWorld, 64); memcpy(m.w2o, worldToObject, 64);
	// Cylinder::getAABB (cylinder.cpp:186-205): the two end circles, component by component
	{
		const float *M = m.o2w;
		float lo[3], hi[3];
		for (int i = 0; i < 3; ++i) {
			volatile float x1 = M[4 * i] * radius, x2 = M[4 * i + 1] * radius;      // objectToWorld(Vector(r,0,0))[i], (0,r,0)[i]
			volatile float p0 = M[4 * i + 3];                                       // objectToWorld(Point(0,0,0))[i] (affine)
			volatile float p1 = M[4 * i + 2] * length; p1 = p1 + M[4 * i + 3];      // objectToWorld(Point(0,0,length))[i]
			volatile float a = x1 * x1, b = x2 * x2; volatile float sum = a + b;
			const float range = std::sqrt((float) sum);
			volatile float l0 = p0 - range, l1 = p1 - range, h0 = p0 + range, h1 = p1 + range;
			lo[i] = std::min((float) l0, (float) l1); hi[i] = std::max((float) h0, (float) h1);
		}
		for (int k = 0; k < 3; ++k) { m.pos[k] = lo[k]; m.pos[3 + k] = hi[k]; }
	}
	ctx->meshes.push_back(std::move(m));
	ctx->built = false;
	return MTSGPU_OK;
}

int mtsgpu_upload_group_mesh(mtsgpu_ctx *ctx, int groupId, int meshIndex, const float *positions, const float *normals,
		const float *texcoords, uint32_t nVert, const uint32_t *triangles, uint32_t nTri, int materialId) {
	ENTER(ctx);
	REQUIRE(ctx, groupId >= 0 && (size_t) groupId <= ctx->groups.size(), MTSGPU_ERR_ARG,
		"upload_group_mesh: groups must be created in order (got %d, expected <= %zu)", groupId, ctx->groups.size());
	if ((size_t) groupId == ctx->groups.size()) {
		mtsgpu_ctx *g = nullptr;
		REQUIRE(ctx, mtsgpu_create(ctx->device, &g) == MTSGPU_OK, MTSGPU_ERR_CUDA, "upload_group_mesh: %s", mtsgpu_last_error(nullptr));
		ctx->groups.push_back(g);
	}
	mtsgpu_ctx *g = ctx->groups[groupId];
	const int rc = mtsgpu_upload_mesh(g, meshIndex, positions, normals, texcoords, nVert, triangles, nTri, materialId, -1);
	if (rc != MTSGPU_OK) return setError(ctx, rc, "group %d: %s", groupId, g->error.c_str());
	ctx->built = false;
	return MTSGPU_OK;
}

int mtsgpu_upload_instance(mtsgpu_ctx *ctx, int shapeIndex, int groupId, const float objectToWorld[16], const float worldToObject[16]) {
	ENTER(ctx);
	REQUIRE(ctx, objectToWorld && worldToObject, MTSGPU_ERR_ARG, "upload_instance: null argument");
	REQUIRE(ctx, shapeIndex >= 0 && (size_t) shapeIndex <= ctx->meshes.size(), MTSGPU_ERR_ARG,
		"upload_instance: shapes must be uploaded in kd-tree order (got %d, expected <= %zu)", shapeIndex, ctx->meshes.size());
	// "A reference to a 'shapegroup' must be specified!" (instance.cpp:42-45)
	REQUIRE(ctx, groupId >= 0 && (size_t) groupId < ctx->groups.size(), MTSGPU_ERR_ARG, "upload_instance: shape group %d was not uploaded", groupId);
	if ((size_t) shapeIndex < ctx->meshes.size()) ctx->meshes.resize(shapeIndex);
	HostMesh m;
	m.nVert = 2; m.nTri = 1; m.materialId = -1; m.emitterId = -1; m.hasNormals = false;
	m.pos.assign(6, 0.0f);
	m.tri = { 0u, INSTANCE_MARK, 1u };
	m.instance = true; m.groupId = groupId;
	memcpy(m.o2w, objectToWorld, 64); memcpy(m.w2o, worldToObject, 64);
	ctx->meshes.push_back(std::move(m));
	ctx->built = false;
	return MTSGPU_OK;
}

int mtsgpu_set_camera(mtsgpu_ctx *ctx, const float rasterToCamera[16], const float cameraToWorld[16], float nearClip, float farClip,
		float apertureRadius, float focusDepth, float shutterOpen, float shutterOpenTime) {
	ENTER(ctx);
	REQUIRE(ctx, rasterToCamera && cameraToWorld, MTSGPU_ERR_ARG, "set_camera: null matrix");
	memcpy(ctx->camera.rasterToCamera, rasterToCamera, 64);
	memcpy(ctx->camera.cameraToWorld, cameraToWorld, 64);
	ctx->camera.nearClip = nearClip; ctx->camera.farClip = farClip;
	ctx->camera.apertureRadius = apertureRadius; ctx->camera.focusDepth = focusDepth;
	ctx->camera.shutterOpen = shutterOpen; ctx->camera.shutterOpenTime = shutterOpenTime;
	ctx->camera.orthographic = 0;
	ctx->haveCamera = true;
	ctx->lumImportance = -1;          // the background luminaire's sphere depends on the camera position
	return MTSGPU_OK;
}

int mtsgpu_set_camera_orthographic(mtsgpu_ctx *ctx, const float rasterToCamera[16], const float cameraToWorld[16], float nearClip, float farClip,
		float shutterOpen, float shutterOpenTime) {
	const int rc = mtsgpu_set_camera(ctx, rasterToCamera, cameraToWorld, nearClip, farClip, 0.0f, 0.0f, shutterOpen, shutterOpenTime);
	if (rc == MTSGPU_OK) ctx->camera.orthographic = 1;
	return rc;
}

int mtsgpu_set_film(mtsgpu_ctx *ctx, int width, int height, int cropX, int cropY, int cropW, int cropH,
		const float filterTable[256], float filterSizeX, float filterSizeY) {
	ENTER(ctx);
	REQUIRE(ctx, filterTable, MTSGPU_ERR_ARG, "set_film: null filter table");
	// same validity rule as Film::Film (src/librender/film.cpp:40-45)
	REQUIRE(ctx, !(cropX < 0 || cropY < 0 || cropW <= 0 || cropH <= 0 || cropX + cropW > width || cropY + cropH > height),
		MTSGPU_ERR_ARG, "Invalid crop window specification!");
	REQUIRE(ctx, filterSizeX > 0 && filterSizeY > 0, MTSGPU_ERR_ARG, "set_film: filter size must be positive");
	FilmDev &f = ctx->film;
	f.width = width; f.height = height; f.cropX = cropX; f.cropY = cropY; f.cropW = cropW; f.cropH = cropH;
	f.filterSizeX = filterSizeX; f.filterSizeY = filterSizeY;
	f.factorX = 15.0f / filterSizeX; f.factorY = 15.0f / filterSizeY;      // FILTER_RESOLUTION / m_size (rfilter.cpp:43-45)
	f.border = (int) std::ceil(std::max(filterSizeX, filterSizeY) - 0.5f);  // renderproc.cpp:142-144
	REQUIRE(ctx, f.border <= 8, MTSGPU_ERR_LIMIT, "set_film: filter radius %g exceeds the supported maximum", std::max(filterSizeX, filterSizeY));
	memcpy(ctx->filterTable, filterTable, sizeof(ctx->filterTable));
	ctx->haveFilm = true;
	return MTSGPU_OK;
}

int mtsgpu_build_accel(mtsgpu_ctx *ctx) {
	ENTER(ctx);
	// ---- shape groups first (ShapeGroup::configure builds its kd-tree, shapegroup.cpp:44-47): each one is a context of
	// its own that sees the scene's materials and textures, so that material ids mean the same thing in both
	bool anyInstance = false;
	for (const HostMesh &m : ctx->meshes) anyInstance |= m.instance;
	if (anyInstance) {
		for (size_t g = 0; g < ctx->groups.size(); ++g) {
			mtsgpu_ctx *c = ctx->groups[g];
			if (c->built) continue;
			c->materials = ctx->materials; c->textures = ctx->textures; c->bitmaps = ctx->bitmaps;
			c->emitters.clear();
			const int rc = buildAccel(c);
			if (rc != MTSGPU_OK) return setError(ctx, rc, "shape group %zu: %s", g, c->error.c_str());
		}
		for (HostMesh &m : ctx->meshes) {
			if (!m.instance) continue;
			// Instance::getAABB (instance.cpp:47-56): the eight corners of the group's kd-tree box through objectToWorld
			const DeviceScene &gs = ctx->groups[m.groupId]->scene;
			float lo[3] = { INFINITY, INFINITY, INFINITY }, hi[3] = { -INFINITY, -INFINITY, -INFINITY };
			for (int corner = 0; corner < 8; ++corner) {
				const float p[3] = { (corner & 1) ? gs.aabbMax[0] : gs.aabbMin[0], (corner & 2) ? gs.aabbMax[1] : gs.aabbMin[1],
					(corner & 4) ? gs.aabbMax[2] : gs.aabbMin[2] };
				float q[4];
				for (int r = 0; r < 4; ++r) {          // Transform::operator()(Point) (transform.h:100-115)
					volatile float a = m.o2w[4 * r] * p[0]; volatile float b = m.o2w[4 * r + 1] * p[1]; volatile float c2 = m.o2w[4 * r + 2] * p[2];
					volatile float sum = a + b; sum = sum + c2; sum = sum + m.o2w[4 * r + 3];
					q[r] = sum;
				}
				if (q[3] != 1.0f) { const float inv = 1.0f / q[3]; for (int k = 0; k < 3; ++k) q[k] *= inv; }
				for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], q[k]); hi[k] = std::max(hi[k], q[k]); }
			}
			for (int k = 0; k < 3; ++k) { m.pos[k] = lo[k]; m.pos[3 + k] = hi[k]; }
		}
	}
	return buildAccel(ctx);
}

int mtsgpu_accel_info_get(mtsgpu_ctx *ctx, mtsgpu_accel_info *info) {
	ENTER(ctx);
	REQUIRE(ctx, ctx->built && info, MTSGPU_ERR_ARG, "accel_info: call mtsgpu_build_accel first");
	*info = ctx->accel;
	return MTSGPU_OK;
}

int mtsgpu_set_progress(mtsgpu_ctx *ctx, mtsgpu_progress_fn fn, void *user, double minIntervalSeconds, int wantFilm) {
	ENTER(ctx);
	REQUIRE(ctx, minIntervalSeconds >= 0, MTSGPU_ERR_ARG, "set_progress: negative interval");
	ctx->progressFn = fn; ctx->progressUser = user; ctx->progressInterval = minIntervalSeconds; ctx->progressWantFilm = fn && wantFilm;
	return MTSGPU_OK;
}

int mtsgpu_render_device(mtsgpu_ctx *ctx, const mtsgpu_render_params *params, volatile const int *cancel, float *d_outFilm, mtsgpu_stats *stats) {
	ENTER(ctx);
	REQUIRE(ctx, params && d_outFilm, MTSGPU_ERR_ARG, "render: null argument");
	return renderDevice(ctx, params, cancel, d_outFilm, stats);
}

int mtsgpu_render(mtsgpu_ctx *ctx, const mtsgpu_render_params *params, volatile const int *cancel, float *outFilm, mtsgpu_stats *stats) {
	ENTER(ctx);
	REQUIRE(ctx, params && outFilm, MTSGPU_ERR_ARG, "render: null argument");
	REQUIRE(ctx, ctx->haveFilm, MTSGPU_ERR_ARG, "render: call mtsgpu_set_film first");
	const size_t bytes = (size_t) ctx->film.cropW * ctx->film.cropH * 5 * sizeof(float);
	CUDA_TRY(ctx, ctx->dFilm.alloc(bytes));
	int rc = renderDevice(ctx, params, cancel, ctx->dFilm.as<float>(), stats);
	if (rc) return rc;
	CUDA_TRY(ctx, cudaMemcpyAsync(outFilm, ctx->dFilm.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
	CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	return MTSGPU_OK;
}

int mtsgpu_trace_device(mtsgpu_ctx *ctx, const float *d_rays, size_t n, int mode, void *d_out, float *elapsedMs) {
	ENTER(ctx);
	REQUIRE(ctx, n == 0 || (d_rays && d_out), MTSGPU_ERR_ARG, "trace: null buffer");
	uint64_t *cnt = nullptr;
	if (ctx->countersEnabled) {
		CUDA_TRY(ctx, ctx->dCounters.alloc(64));
		CUDA_TRY(ctx, cudaMemsetAsync(ctx->dCounters.p, 0, 16, ctx->stream));
		cnt = ctx->dCounters.as<uint64_t>();
	}
	CUDA_TRY(ctx, cudaEventRecord(ctx->ev[4], ctx->stream));
	int rc = launchTrace(ctx, reinterpret_cast<const float4 *>(d_rays), n, mode, d_out, cnt);
	if (rc) return rc;
	CUDA_TRY(ctx, cudaEventRecord(ctx->ev[5], ctx->stream));
	CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	CUDA_TRY(ctx, cudaGetLastError());
	if (elapsedMs) { float ms = 0; cudaEventElapsedTime(&ms, ctx->ev[4], ctx->ev[5]); *elapsedMs = ms; }
	{ const int frc = checkTraceFlags(ctx); if (frc) return frc; }
	if (cnt) {
		uint64_t h[2];
		CUDA_TRY(ctx, cudaMemcpy(h, cnt, 16, cudaMemcpyDeviceToHost));
		ctx->lastCounters.rays = n; ctx->lastCounters.nodeVisits = h[0]; ctx->lastCounters.triTests = h[1];
	}
	return MTSGPU_OK;
}

int mtsgpu_trace(mtsgpu_ctx *ctx, const float *rays, size_t n, int mode, void *out) {
	ENTER(ctx);
	REQUIRE(ctx, n == 0 || (rays && out), MTSGPU_ERR_ARG, "trace: null buffer");
	REQUIRE(ctx, mode == MTSGPU_TRACE_CLOSEST || mode == MTSGPU_TRACE_ANY, MTSGPU_ERR_ARG, "trace: unknown mode %d", mode);
	if (n == 0) return MTSGPU_OK;
	const size_t outBytes = n * (mode == MTSGPU_TRACE_CLOSEST ? sizeof(mtsgpu_hit) : sizeof(uint32_t));
	CUDA_TRY(ctx, ctx->dRays.alloc(n * 32));
	CUDA_TRY(ctx, ctx->dHits.alloc(outBytes));
	const size_t outStride = mode == MTSGPU_TRACE_CLOSEST ? sizeof(mtsgpu_hit) : sizeof(uint32_t);
	if (n < (1u << 21) || ctx->countersEnabled) {
		CUDA_TRY(ctx, cudaMemcpyAsync(ctx->dRays.p, rays, n * 32, cudaMemcpyHostToDevice, ctx->stream));
		int rc = mtsgpu_trace_device(ctx, ctx->dRays.as<float>(), n, mode, ctx->dHits.p, nullptr);
		if (rc) return rc;
		CUDA_TRY(ctx, cudaMemcpyAsync(out, ctx->dHits.p, outBytes, cudaMemcpyDeviceToHost, ctx->stream));
		CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
		return MTSGPU_OK;
	}
	// Large batches: three-stage pipeline over chunks — upload of chunk k+1, traversal of chunk k and download of
	// chunk k-1 run concurrently on three streams (PCIe is full duplex), ordered by events. With pinned host
	// buffers the copies are truly asynchronous; with pageable ones CUDA stages them and the result is the same.
	if (!ctx->copyIn) {
		CUDA_TRY(ctx, cudaStreamCreateWithFlags(&ctx->copyIn, cudaStreamNonBlocking));
		CUDA_TRY(ctx, cudaStreamCreateWithFlags(&ctx->copyOut, cudaStreamNonBlocking));
		for (int i = 0; i < 2 * MTSGPU_PIPE_CHUNKS; ++i) CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->pipeEv[i], cudaEventDisableTiming));
	}
	const size_t chunk = std::max<size_t>((size_t) 1 << 20, (n + MTSGPU_PIPE_CHUNKS - 1) / MTSGPU_PIPE_CHUNKS);
	int k = 0;
	for (size_t off = 0; off < n; off += chunk, ++k) {
		const size_t m = std::min(chunk, n - off);
		float *dIn = ctx->dRays.as<float>() + off * 8;
		char *dOut = ctx->dHits.as<char>() + off * outStride;
		CUDA_TRY(ctx, cudaMemcpyAsync(dIn, rays + off * 8, m * 32, cudaMemcpyHostToDevice, ctx->copyIn));
		CUDA_TRY(ctx, cudaEventRecord(ctx->pipeEv[2 * k], ctx->copyIn));
		CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, ctx->pipeEv[2 * k], 0));
		int rc = launchTrace(ctx, reinterpret_cast<const float4 *>(dIn), m, mode, dOut, nullptr);
		if (rc) {      // no copy may still be in flight into the caller's buffers when this call returns
			cudaStreamSynchronize(ctx->copyIn); cudaStreamSynchronize(ctx->stream); cudaStreamSynchronize(ctx->copyOut);
			return rc;
		}
		CUDA_TRY(ctx, cudaEventRecord(ctx->pipeEv[2 * k + 1], ctx->stream));
		CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->copyOut, ctx->pipeEv[2 * k + 1], 0));
		CUDA_TRY(ctx, cudaMemcpyAsync(static_cast<char *>(out) + off * outStride, dOut, m * outStride, cudaMemcpyDeviceToHost, ctx->copyOut));
	}
	CUDA_TRY(ctx, cudaStreamSynchronize(ctx->copyOut));
	CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
	CUDA_TRY(ctx, cudaGetLastError());
	return checkTraceFlags(ctx);
}

int mtsgpu_bsdf_eval(mtsgpu_ctx *ctx, const float *queries, size_t n, float *out) { ENTER(ctx); return hookBsdfEval(ctx, queries, n, out); }
int mtsgpu_sampler_values(mtsgpu_ctx *ctx, const mtsgpu_render_params *params, uint32_t pixel, uint32_t draw, float *out) {
	ENTER(ctx);
	REQUIRE(ctx, params && out, MTSGPU_ERR_ARG, "sampler_values: null argument");
	return hookSamplerValues(ctx, params, pixel, draw, out);
}
int mtsgpu_intersect_uv(mtsgpu_ctx *ctx, const float *rays, size_t n, float *out) { ENTER(ctx); return hookIntersectUv(ctx, rays, n, out); }
int mtsgpu_texture_lookup(mtsgpu_ctx *ctx, const float *samples, size_t n, int sampleCount, float *out) { ENTER(ctx); return hookTextureLookup(ctx, samples, n, sampleCount, out); }
int mtsgpu_emitter_sample(mtsgpu_ctx *ctx, const float *queries, size_t n, float *out) { ENTER(ctx); return hookEmitterSample(ctx, queries, n, out); }
int mtsgpu_camera_rays(mtsgpu_ctx *ctx, const float *samples, size_t n, float *out) { ENTER(ctx); return hookCameraRays(ctx, samples, n, out); }
int mtsgpu_film_splat(mtsgpu_ctx *ctx, const float *samples, size_t n, float *outFilm) { ENTER(ctx); return hookFilmSplat(ctx, samples, n, outFilm); }
int mtsgpu_intersect_full(mtsgpu_ctx *ctx, const float *rays, size_t n, mtsgpu_hit *hits, float *out) { ENTER(ctx); return hookIntersectFull(ctx, rays, n, hits, out); }

int mtsgpu_get_triaccel(mtsgpu_ctx *ctx, void *outRecords, size_t maxRecords, float aabb6[6]) {
	ENTER(ctx);
	REQUIRE(ctx, ctx->built, MTSGPU_ERR_ARG, "get_triaccel: call mtsgpu_build_accel first");
	const size_t n = std::min<size_t>(maxRecords, ctx->accel.triangles);
	if (outRecords && n)
		CUDA_TRY(ctx, cudaMemcpy(outRecords, ctx->dTriAccelOrig.p, n * sizeof(TriAccelRec), cudaMemcpyDeviceToHost));
	if (aabb6) {
		for (int c = 0; c < 3; ++c) { aabb6[c] = ctx->scene.aabbMin[c]; aabb6[3 + c] = ctx->scene.aabbMax[c]; }
	}
	return MTSGPU_OK;
}

int mtsgpu_set_counters(mtsgpu_ctx *ctx, int enabled) {
	ENTER(ctx);
	if (enabled && !ctx->countersEnabled && ctx->dTraceFlags.p)      // word 1 of the flag buffer: deepest stack seen by the counter builds
		CUDA_TRY(ctx, cudaMemset((char *) ctx->dTraceFlags.p + 4, 0, 4));
	ctx->countersEnabled = enabled != 0;
	return MTSGPU_OK;
}

int mtsgpu_get_counters(mtsgpu_ctx *ctx, mtsgpu_trace_counters *out) {
	ENTER(ctx);
	REQUIRE(ctx, out, MTSGPU_ERR_ARG, "get_counters: null argument");
	*out = ctx->lastCounters;
	if (ctx->dTraceFlags.p) {
		uint32_t deepest = 0;
		CUDA_TRY(ctx, cudaMemcpy(&deepest, (const char *) ctx->dTraceFlags.p + 4, 4, cudaMemcpyDeviceToHost));
		out->maxStackEntries = deepest;
	}
	return MTSGPU_OK;
}

} // extern "C"

/* TEST INFRASTRUCTURE — reader for the little-endian scene interchange file the Python
 * scene generators write (mitsuba-renderer_b200/scenefile.py documents the same layout).
 * Used only by the oracle harness; the CUDA library takes plain arrays through its C ABI. */
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include <stdexcept>

namespace sio {

enum MaterialType { MAT_LAMBERTIAN = 0, MAT_MICROFACET = 1, MAT_MIRROR = 2, MAT_ROUGHMETAL = 3, MAT_DIELECTRIC = 4 };

struct Material { uint32_t type; float p[16]; uint32_t tex[3]; };
struct Texture  { uint32_t type; float p[12]; float gamma; uint32_t width, height; std::vector<uint8_t> pixels; /* bitmap: RGB, top row first */ std::string exrFile; /* exrtexture instead */ };
struct Emitter  { float intensity[3]; float samplingWeight; uint32_t type; float p[11]; float aux[16]; std::string envFile; /* type 5 */ };
struct Shape {
	uint32_t kind;   /* 0 triangle mesh, 1 analytic sphere, 2 instance of a shape group, 3 analytic cylinder (toWorld) */
	uint32_t groupId; float toWorld[16];
	float center[3], radius; uint32_t inverted;
	uint32_t nVert, nTri; int32_t matId, emitId; uint32_t hasNormals, hasUv;
	std::vector<float> pos, nrm, uv; std::vector<uint32_t> tri;
};

struct Scene {
	float cameraToWorld[16];
	float fov, nearClip, farClip, apertureRadius, focusDepth;
	uint32_t mapSmallerSide;
	uint32_t width, height, cropX, cropY, cropW, cropH, filterType;
	uint32_t spp, mode; int32_t maxDepth, rrDepth; uint32_t strictNormals, nLumSamples, nBsdfSamples;
	float shutterOpen, shutterClose; uint32_t importanceSampleLuminaires;
	uint32_t samplerType, samplerDepth;
	std::vector<Texture> textures;
	std::vector<Material> materials;
	std::vector<Emitter> emitters;
	std::vector<Shape> shapes;
	std::vector<std::vector<Shape> > groups;   /* shape groups: meshes in object space */
};

struct Reader {
	FILE *f;
	explicit Reader(const std::string &path) : f(fopen(path.c_str(), "rb")) {
		if (!f) throw std::runtime_error("cannot open " + path);
	}
	~Reader() { if (f) fclose(f); }
	void raw(void *dst, size_t n) { if (n && fread(dst, 1, n, f) != n) throw std::runtime_error("short read"); }
	template <typename T> T get() { T v; raw(&v, sizeof(T)); return v; }
};

inline void load(const std::string &path, Scene &s) {
	Reader r(path);
	char magic[8]; r.raw(magic, 8);
	if (memcmp(magic, "MTSB200", 8) != 0) throw std::runtime_error("bad magic in " + path);
	if (r.get<uint32_t>() != 11) throw std::runtime_error("bad version in " + path);
	r.raw(s.cameraToWorld, 64);
	s.fov = r.get<float>(); s.nearClip = r.get<float>(); s.farClip = r.get<float>();
	s.apertureRadius = r.get<float>(); s.focusDepth = r.get<float>();
	s.mapSmallerSide = r.get<uint32_t>();
	s.width = r.get<uint32_t>(); s.height = r.get<uint32_t>();
	s.cropX = r.get<uint32_t>(); s.cropY = r.get<uint32_t>();
	s.cropW = r.get<uint32_t>(); s.cropH = r.get<uint32_t>();
	s.filterType = r.get<uint32_t>();
	s.spp = r.get<uint32_t>(); s.mode = r.get<uint32_t>();
	s.maxDepth = r.get<int32_t>(); s.rrDepth = r.get<int32_t>();
	s.strictNormals = r.get<uint32_t>(); s.nLumSamples = r.get<uint32_t>(); s.nBsdfSamples = r.get<uint32_t>();
	s.samplerType = r.get<uint32_t>(); s.samplerDepth = r.get<uint32_t>();
	s.shutterOpen = r.get<float>(); s.shutterClose = r.get<float>(); s.importanceSampleLuminaires = r.get<uint32_t>();
	s.textures.resize(r.get<uint32_t>());
	for (auto &t : s.textures) {
		t.type = r.get<uint32_t>(); r.raw(t.p, 48);
		t.gamma = -1.0f; t.width = t.height = 0;
		if (t.type == 1 && r.get<uint32_t>() == 1) {
			t.exrFile.resize(r.get<uint32_t>()); r.raw(&t.exrFile[0], t.exrFile.size());
		} else if (t.type == 1) {
			t.gamma = r.get<float>(); t.width = r.get<uint32_t>(); t.height = r.get<uint32_t>();
			t.pixels.resize((size_t) t.width * t.height * 3); r.raw(t.pixels.data(), t.pixels.size());
		}
	}
	s.materials.resize(r.get<uint32_t>());
	for (auto &m : s.materials) { m.type = r.get<uint32_t>(); r.raw(m.p, 64); r.raw(m.tex, 12); }
	s.emitters.resize(r.get<uint32_t>());
	for (auto &e : s.emitters) { r.raw(e.intensity, 12); e.samplingWeight = r.get<float>(); e.type = r.get<uint32_t>(); r.raw(e.p, 44); r.raw(e.aux, 64);
		if (e.type == 5) { e.envFile.resize(r.get<uint32_t>()); r.raw(&e.envFile[0], e.envFile.size()); } }
	auto readMesh = [&](Shape &sh) {
		sh.kind = 0;
		sh.nVert = r.get<uint32_t>(); sh.nTri = r.get<uint32_t>();
		sh.matId = r.get<int32_t>(); sh.emitId = r.get<int32_t>();
		sh.hasNormals = r.get<uint32_t>(); sh.hasUv = r.get<uint32_t>();
		sh.pos.resize(3 * (size_t) sh.nVert); r.raw(sh.pos.data(), sh.pos.size() * 4);
		if (sh.hasNormals) { sh.nrm.resize(3 * (size_t) sh.nVert); r.raw(sh.nrm.data(), sh.nrm.size() * 4); }
		if (sh.hasUv) { sh.uv.resize(2 * (size_t) sh.nVert); r.raw(sh.uv.data(), sh.uv.size() * 4); }
		sh.tri.resize(3 * (size_t) sh.nTri); r.raw(sh.tri.data(), sh.tri.size() * 4);
	};
	s.groups.resize(r.get<uint32_t>());
	for (auto &g : s.groups) { g.resize(r.get<uint32_t>()); for (auto &sh : g) readMesh(sh); }
	s.shapes.resize(r.get<uint32_t>());
	for (auto &sh : s.shapes) {
		sh.kind = r.get<uint32_t>();
		if (sh.kind == 3) {
			sh.nVert = sh.nTri = sh.hasNormals = sh.hasUv = 0; sh.emitId = -1;
			sh.matId = r.get<int32_t>(); r.raw(sh.toWorld, 64);
			continue;
		}
		if (sh.kind == 2) {
			sh.nVert = sh.nTri = sh.hasNormals = sh.hasUv = 0; sh.matId = sh.emitId = -1;
			sh.groupId = r.get<uint32_t>(); r.raw(sh.toWorld, 64);
			continue;
		}
		if (sh.kind == 1) {
			sh.nVert = sh.nTri = sh.hasNormals = sh.hasUv = 0;
			sh.matId = r.get<int32_t>(); sh.emitId = r.get<int32_t>();
			r.raw(sh.center, 12); sh.radius = r.get<float>(); sh.inverted = r.get<uint32_t>();
			continue;
		}
		sh.nVert = r.get<uint32_t>(); sh.nTri = r.get<uint32_t>();
		sh.matId = r.get<int32_t>(); sh.emitId = r.get<int32_t>();
		sh.hasNormals = r.get<uint32_t>(); sh.hasUv = r.get<uint32_t>();
		sh.pos.resize(3 * (size_t) sh.nVert); r.raw(sh.pos.data(), sh.pos.size() * 4);
		if (sh.hasNormals) { sh.nrm.resize(3 * (size_t) sh.nVert); r.raw(sh.nrm.data(), sh.nrm.size() * 4); }
		if (sh.hasUv) { sh.uv.resize(2 * (size_t) sh.nVert); r.raw(sh.uv.data(), sh.uv.size() * 4); }
		sh.tri.resize(3 * (size_t) sh.nTri); r.raw(sh.tri.data(), sh.tri.size() * 4);
	}
}

inline std::vector<float> loadFloats(const std::string &path) {
	Reader r(path);
	fseek(r.f, 0, SEEK_END); long n = ftell(r.f); fseek(r.f, 0, SEEK_SET);
	std::vector<float> v((size_t) n / 4);
	r.raw(v.data(), v.size() * 4);
	return v;
}

inline void save(const std::string &path, const void *data, size_t bytes) {
	FILE *f = fopen(path.c_str(), "wb");
	if (!f) throw std::runtime_error("cannot write " + path);
	if (bytes && fwrite(data, 1, bytes, f) != bytes) { fclose(f); throw std::runtime_error("short write"); }
	fclose(f);
}

} // namespace sio

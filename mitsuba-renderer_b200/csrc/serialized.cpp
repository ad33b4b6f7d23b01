// `.serialized` triangle-mesh files: the on-disk format either side of the hot path.
// Host code only (zlib + plain C++). Restates, from the format's definition in the reference,
//   TriMesh::TriMesh(Stream *, int index)        src/librender/trimesh.cpp:156-237   (reader)
//   TriMesh::serialize(Stream *)                 src/librender/trimesh.cpp:748-790   (writer)
//   the trailing offset table                    src/converter/converter.cpp:293-296
//   SerializedMesh::SerializedMesh               src/shapes/serialized.cpp:30-91     (toWorld, winding)
//   TriMesh::computeNormals                      src/librender/trimesh.cpp:472-540   (vertex normals)
// so that a mesh loaded here is, value for value, the mesh the reference's `serialized` shape would
// hand to its kd-tree: it then goes to the GPU through the same path as mtsgpu_upload_mesh.
#include "context.h"
#include <zlib.h>
#include <cstring>
#include <cstdlib>
#include <cmath>
#include <cstdio>
#include <vector>
#include <string>
#include <algorithm>

void setGlobalError(const char *fmt, ...);      // api.cu: message behind mtsgpu_last_error(NULL)

namespace {

const uint16_t FILEFORMAT_HEADER = 0x041C, FILEFORMAT_VERSION_V3 = 0x0003;     // trimesh.cpp:30-31
enum { EHasNormals = 0x0001, EHasTexcoords = 0x0002, EHasColors = 0x0008, EFaceNormals = 0x0010,
       ESinglePrecision = 0x1000, EDoublePrecision = 0x2000 };                // trimesh.cpp:72-80

bool readFile(const char *path, std::vector<unsigned char> &data) {
	FILE *f = fopen(path, "rb");
	if (!f) { setGlobalError("serialized: cannot open \"%s\"", path); return false; }
	fseek(f, 0, SEEK_END); const long size = ftell(f); fseek(f, 0, SEEK_SET);
	data.resize(size > 0 ? (size_t) size : 0);
	const size_t got = data.empty() ? 0 : fread(data.data(), 1, data.size(), f);
	fclose(f);
	if (got != data.size()) { setGlobalError("serialized: short read on \"%s\"", path); return false; }
	return true;
}

inline uint32_t le32(const unsigned char *p) { return (uint32_t) p[0] | ((uint32_t) p[1] << 8) | ((uint32_t) p[2] << 16) | ((uint32_t) p[3] << 24); }
inline uint16_t le16(const unsigned char *p) { return (uint16_t) (p[0] | (p[1] << 8)); }

// Inflating reader over one zlib stream (what ZStream does for reads, zstream.cpp)
struct Inflater {
	z_stream zs; bool ok;
	Inflater(const unsigned char *src, size_t n) {
		memset(&zs, 0, sizeof(zs));
		ok = inflateInit(&zs) == Z_OK;
		zs.next_in = const_cast<unsigned char *>(src);
		zs.avail_in = (uInt) std::min<size_t>(n, 0xFFFFFFFFu);
	}
	~Inflater() { if (ok) inflateEnd(&zs); }
	bool read(void *dst, size_t n) {
		unsigned char *out = static_cast<unsigned char *>(dst);
		while (n > 0) {
			const uInt chunk = (uInt) std::min<size_t>(n, 1u << 30);
			zs.next_out = out; zs.avail_out = chunk;
			const int rc = inflate(&zs, Z_NO_FLUSH);
			const size_t produced = chunk - zs.avail_out;
			out += produced; n -= produced;
			if (rc == Z_STREAM_END && n > 0) return false;
			if (rc != Z_OK && rc != Z_STREAM_END) return false;
			if (produced == 0 && zs.avail_in == 0) return false;
		}
		return true;
	}
};

// readHelper (trimesh.cpp:135-154): floats are stored in the file's precision
bool readFloats(Inflater &in, bool fileDouble, float *dst, size_t count) {
	if (!fileDouble) return in.read(dst, count * sizeof(float));
	std::vector<double> tmp(count);
	if (!in.read(tmp.data(), count * sizeof(double))) return false;
	for (size_t i = 0; i < count; ++i) dst[i] = (float) tmp[i];
	return true;
}

template <typename T> T *allocArray(size_t n) { return static_cast<T *>(malloc(std::max<size_t>(n, 1) * sizeof(T))); }

struct V { float x, y, z; };
inline V sub(V a, V b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
inline V add(V a, V b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
inline V mul(V a, float s) { return { a.x * s, a.y * s, a.z * s }; }
inline float dot(V a, V b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float len(V a) { return std::sqrt(dot(a, a)); }
inline V crossv(V a, V b) { return { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }
inline V divv(V a, float f) { const float r = 1.0f / f; return mul(a, r); }                 // TVector3::operator/ (vector.h:311-318)
inline V normalizev(V a) { return divv(a, len(a)); }
inline float unitAngle(V u, V v) {                                                            // core/util.h:343-348
	if (dot(u, v) < 0) return 3.14159265358979323846f - 2 * std::asin(len(add(v, u)) / 2);
	return 2 * std::asin(len(sub(v, u)) / 2);
}

// TriMesh::computeNormals (trimesh.cpp:472-540)
void computeNormals(mtsgpu_mesh_data &m, bool faceNormals, bool flipNormals) {
	if (faceNormals) {
		free(m.normals); m.normals = nullptr;
		if (flipNormals)
			for (uint32_t i = 0; i < m.nTri; ++i) std::swap(m.triangles[3 * i], m.triangles[3 * i + 1]);
		return;
	}
	if (m.normals) {
		if (flipNormals)
			for (size_t i = 0; i < 3 * (size_t) m.nVert; ++i) m.normals[i] *= -1;
		return;
	}
	m.normals = allocArray<float>(3 * (size_t) m.nVert);
	memset(m.normals, 0, 3 * (size_t) m.nVert * sizeof(float));
	const V *P = reinterpret_cast<const V *>(m.positions);
	V *N = reinterpret_cast<V *>(m.normals);
	// "Computing Vertex Normals from Polygonal Facets" (Thuermer & Wuethrich), accumulated in triangle order
	for (uint32_t t = 0; t < m.nTri; ++t) {
		const uint32_t *idx = m.triangles + 3 * (size_t) t;
		V n = { 0, 0, 0 };
		for (int i = 0; i < 3; ++i) {
			const V v0 = P[idx[i]], v1 = P[idx[(i + 1) % 3]], v2 = P[idx[(i + 2) % 3]];
			const V sideA = sub(v1, v0), sideB = sub(v2, v0);
			if (i == 0) {
				n = crossv(sideA, sideB);
				const float length = len(n);
				if (length == 0) break;
				n = divv(n, length);
			}
			const float angle = unitAngle(normalizev(sideA), normalizev(sideB));
			N[idx[i]] = add(N[idx[i]], mul(n, angle));
		}
	}
	for (uint32_t i = 0; i < m.nVert; ++i) {
		float length = len(N[i]);
		if (flipNormals) length *= -1;
		if (length != 0) N[i] = divv(N[i], length);
		else N[i] = { 1, 0, 0 };
	}
}

bool isIdentity(const float *m) {
	if (!m) return true;
	for (int i = 0; i < 16; ++i) if (m[i] != ((i % 5 == 0) ? 1.0f : 0.0f)) return false;
	return true;
}

// Gauss-Jordan inverse in double for callers that pass no inverse (the reference keeps its own
// float inverse next to the matrix, transform.h:40-60)
bool invert4(const float *m, float *out) {
	double a[4][8];
	for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { a[i][j] = m[4 * i + j]; a[i][4 + j] = i == j; }
	for (int c = 0; c < 4; ++c) {
		int piv = c;
		for (int r = c + 1; r < 4; ++r) if (std::fabs(a[r][c]) > std::fabs(a[piv][c])) piv = r;
		if (a[piv][c] == 0) return false;
		for (int j = 0; j < 8; ++j) std::swap(a[c][j], a[piv][j]);
		const double d = a[c][c];
		for (int j = 0; j < 8; ++j) a[c][j] /= d;
		for (int r = 0; r < 4; ++r) if (r != c) { const double f = a[r][c]; for (int j = 0; j < 8; ++j) a[r][j] -= f * a[c][j]; }
	}
	for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out[4 * i + j] = (float) a[i][4 + j];
	return true;
}

} // namespace

extern "C" {

void mtsgpu_serialized_free(mtsgpu_mesh_data *mesh) {
	if (!mesh) return;
	free(mesh->positions); free(mesh->normals); free(mesh->texcoords); free(mesh->colors); free(mesh->triangles);
	memset(mesh, 0, sizeof(*mesh));
}

int mtsgpu_serialized_count(const char *path, int *count) {
	if (!path || !count) return MTSGPU_ERR_ARG;
	std::vector<unsigned char> data;
	if (!readFile(path, data)) return MTSGPU_ERR_ARG;
	if (data.size() < 8 || le16(data.data()) != FILEFORMAT_HEADER) { setGlobalError("serialized: \"%s\": Encountered an invalid file format!", path); return MTSGPU_ERR_ARG; }
	const uint32_t n = le32(data.data() + data.size() - 4);
	// a file written without the table (one mesh, TriMesh::serialize alone) ends inside the zlib stream
	if (n == 0 || (uint64_t) n * 4 + 4 > data.size() || le32(data.data() + data.size() - 4 * ((size_t) n + 1)) != 0) { *count = 1; return MTSGPU_OK; }
	*count = (int) n;
	return MTSGPU_OK;
}

int mtsgpu_serialized_load(const char *path, int shapeIndex, const float *toWorld, const float *toWorldInv,
		int faceNormals, int flipNormals, mtsgpu_mesh_data *out) {
	if (!path || !out) return MTSGPU_ERR_ARG;
	memset(out, 0, sizeof(*out));
	std::vector<unsigned char> data;
	if (!readFile(path, data)) return MTSGPU_ERR_ARG;
	size_t pos = 0;
	if (shapeIndex != 0) {                                           // trimesh.cpp:160-173
		if (data.size() < 4) { setGlobalError("serialized: \"%s\" is truncated", path); return MTSGPU_ERR_ARG; }
		const uint32_t count = le32(data.data() + data.size() - 4);
		if (shapeIndex < 0 || shapeIndex > (int) count || (uint64_t) count * 4 + 4 > data.size()) {
			setGlobalError("Unable to unserialize mesh, shape index is out of range! (requested %i out of 0..%i)", shapeIndex, (int) count - 1);
			return MTSGPU_ERR_ARG;
		}
		pos = le32(data.data() + data.size() - 4 * (1 + (size_t) count - shapeIndex));
	}
	if (pos + 4 > data.size()) { setGlobalError("serialized: \"%s\": bad mesh offset", path); return MTSGPU_ERR_ARG; }
	const uint16_t format = le16(data.data() + pos), version = le16(data.data() + pos + 2);
	if (format == 0x1C04) { setGlobalError("Encountered a geometry file generated by an old version of Mitsuba."); return MTSGPU_ERR_UNSUPPORTED; }
	if (format != FILEFORMAT_HEADER) { setGlobalError("Encountered an invalid file format!"); return MTSGPU_ERR_ARG; }
	if (version != FILEFORMAT_VERSION_V3) { setGlobalError("Encountered an incompatible file version!"); return MTSGPU_ERR_UNSUPPORTED; }
	Inflater in(data.data() + pos + 4, data.size() - pos - 4);
	uint32_t flags = 0; uint64_t nVert = 0, nTri = 0;
	if (!in.ok || !in.read(&flags, 4) || !in.read(&nVert, 8) || !in.read(&nTri, 8) || nVert > 0xFFFFFFFFull || nTri > 0xFFFFFFFFull) {
		setGlobalError("serialized: \"%s\": corrupt mesh header", path); return MTSGPU_ERR_ARG;
	}
	mtsgpu_mesh_data m; memset(&m, 0, sizeof(m));
	m.nVert = (uint32_t) nVert; m.nTri = (uint32_t) nTri; m.flags = flags;
	const bool dbl = (flags & EDoublePrecision) != 0;
	bool ok = true;
	m.positions = allocArray<float>(3 * nVert);
	ok = ok && m.positions && readFloats(in, dbl, m.positions, 3 * nVert);
	if (ok && (flags & EHasNormals)) { m.normals = allocArray<float>(3 * nVert); ok = m.normals && readFloats(in, dbl, m.normals, 3 * nVert); }
	if (ok && (flags & EHasTexcoords)) { m.texcoords = allocArray<float>(2 * nVert); ok = m.texcoords && readFloats(in, dbl, m.texcoords, 2 * nVert); }
	if (ok && (flags & EHasColors)) { m.colors = allocArray<float>(3 * nVert); ok = m.colors && readFloats(in, dbl, m.colors, 3 * nVert); }
	if (ok) { m.triangles = allocArray<uint32_t>(3 * nTri); ok = m.triangles && in.read(m.triangles, 3 * nTri * sizeof(uint32_t)); }
	if (ok)
		for (size_t i = 0; i < 3 * (size_t) nTri; ++i)
			if (m.triangles[i] >= m.nVert) { ok = false; break; }
	if (!ok) { mtsgpu_serialized_free(&m); setGlobalError("serialized: \"%s\": corrupt or truncated mesh data", path); return MTSGPU_ERR_ARG; }

	// ---- SerializedMesh::SerializedMesh (serialized.cpp:70-89): object -> world
	if (!isIdentity(toWorld)) {
		float invBuf[16];
		if (!toWorldInv) {
			if (!invert4(toWorld, invBuf)) { mtsgpu_serialized_free(&m); setGlobalError("serialized: toWorld is singular"); return MTSGPU_ERR_ARG; }
			toWorldInv = invBuf;
		}
		const float *T = toWorld, *I = toWorldInv;
		for (uint32_t i = 0; i < m.nVert; ++i) {                       // Transform::operator()(Point), transform.h:102-119
			float *p = m.positions + 3 * (size_t) i;
			const float x = T[0] * p[0] + T[1] * p[1] + T[2] * p[2] + T[3];
			const float y = T[4] * p[0] + T[5] * p[1] + T[6] * p[2] + T[7];
			const float z = T[8] * p[0] + T[9] * p[1] + T[10] * p[2] + T[11];
			const float w = T[12] * p[0] + T[13] * p[1] + T[14] * p[2] + T[15];
			if (w == 1.0f) { p[0] = x; p[1] = y; p[2] = z; }
			else { const float r = 1.0f / w; p[0] = x * r; p[1] = y * r; p[2] = z * r; }
		}
		if (m.normals)
			for (uint32_t i = 0; i < m.nVert; ++i) {                   // Transform::operator()(Normal): inverse transpose, transform.h:173-181
				float *n = m.normals + 3 * (size_t) i;
				const float x = I[0] * n[0] + I[4] * n[1] + I[8] * n[2];
				const float y = I[1] * n[0] + I[5] * n[1] + I[9] * n[2];
				const float z = I[2] * n[0] + I[6] * n[1] + I[10] * n[2];
				n[0] = x; n[1] = y; n[2] = z;
			}
		const float det = ((T[0] * (T[5] * T[10] - T[6] * T[9])) - (T[1] * (T[4] * T[10] - T[6] * T[8])) + (T[2] * (T[4] * T[9] - T[5] * T[8])));   // matrix.h:617-621
		if (det < 0)
			for (uint32_t i = 0; i < m.nTri; ++i) std::swap(m.triangles[3 * (size_t) i], m.triangles[3 * (size_t) i + 1]);
	}
	computeNormals(m, faceNormals != 0, flipNormals != 0);            // TriMesh::configure (trimesh.cpp:266-286)
	*out = m;
	return MTSGPU_OK;
}

int mtsgpu_serialized_save(const char *path, const mtsgpu_mesh_data *meshes, int count) {
	if (!path || !meshes || count <= 0) return MTSGPU_ERR_ARG;
	std::vector<unsigned char> file;
	std::vector<uint32_t> dict;
	for (int k = 0; k < count; ++k) {
		const mtsgpu_mesh_data &m = meshes[k];
		if (!m.positions || !m.triangles || m.nVert == 0 || m.nTri == 0) { setGlobalError("serialized_save: mesh %d is empty", k); return MTSGPU_ERR_ARG; }
		dict.push_back((uint32_t) file.size());                         // obj.cpp:196: offset of this mesh
		std::vector<unsigned char> raw;
		auto put = [&](const void *p, size_t n) { const unsigned char *b = static_cast<const unsigned char *>(p); raw.insert(raw.end(), b, b + n); };
		uint32_t flags = ESinglePrecision;
		if (m.normals) flags |= EHasNormals;
		if (m.texcoords) flags |= EHasTexcoords;
		if (m.colors) flags |= EHasColors;
		if (m.flags & EFaceNormals) flags |= EFaceNormals;
		const uint64_t nv = m.nVert, nt = m.nTri;
		put(&flags, 4); put(&nv, 8); put(&nt, 8);
		put(m.positions, 12 * (size_t) nv);
		if (m.normals) put(m.normals, 12 * (size_t) nv);
		if (m.texcoords) put(m.texcoords, 8 * (size_t) nv);
		if (m.colors) put(m.colors, 12 * (size_t) nv);
		put(m.triangles, 12 * (size_t) nt);
		uLongf bound = compressBound((uLong) raw.size());
		std::vector<unsigned char> packed(bound);
		if (compress2(packed.data(), &bound, raw.data(), (uLong) raw.size(), Z_DEFAULT_COMPRESSION) != Z_OK) { setGlobalError("serialized_save: deflate failed"); return MTSGPU_ERR_ARG; }
		const unsigned char hdr[4] = { (unsigned char) (FILEFORMAT_HEADER & 0xFF), (unsigned char) (FILEFORMAT_HEADER >> 8), (unsigned char) FILEFORMAT_VERSION_V3, 0 };
		file.insert(file.end(), hdr, hdr + 4);
		file.insert(file.end(), packed.begin(), packed.begin() + bound);
	}
	for (uint32_t off : dict) { const unsigned char *b = reinterpret_cast<const unsigned char *>(&off); file.insert(file.end(), b, b + 4); }   // converter.cpp:293-296
	const uint32_t n = (uint32_t) dict.size();
	{ const unsigned char *b = reinterpret_cast<const unsigned char *>(&n); file.insert(file.end(), b, b + 4); }
	FILE *f = fopen(path, "wb");
	if (!f) { setGlobalError("serialized_save: cannot create \"%s\"", path); return MTSGPU_ERR_ARG; }
	const size_t put = fwrite(file.data(), 1, file.size(), f);
	fclose(f);
	if (put != file.size()) { setGlobalError("serialized_save: short write on \"%s\"", path); return MTSGPU_ERR_ARG; }
	return MTSGPU_OK;
}

int mtsgpu_upload_serialized(mtsgpu_ctx *ctx, int shapeIndex, const char *path, int fileShapeIndex, const float *toWorld,
		const float *toWorldInv, int faceNormals, int flipNormals, int materialId, int emitterId) {
	if (!ctx) return MTSGPU_ERR_ARG;
	mtsgpu_mesh_data m;
	const int rc = mtsgpu_serialized_load(path, fileShapeIndex, toWorld, toWorldInv, faceNormals, flipNormals, &m);
	if (rc != MTSGPU_OK) return setError(ctx, rc, "%s", mtsgpu_last_error(nullptr));
	const int rc2 = mtsgpu_upload_mesh(ctx, shapeIndex, m.positions, m.normals, m.texcoords, m.nVert, m.triangles, m.nTri, materialId, emitterId);
	mtsgpu_serialized_free(&m);
	return rc2;
}

} // extern "C"

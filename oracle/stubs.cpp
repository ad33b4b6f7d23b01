/* TEST INFRASTRUCTURE — link-time stand-ins for the two reference translation units that
 * cannot be compiled here (src/libcore/bitmap.cpp needs OpenEXR/libpng/libjpeg headers,
 * src/librender/testcase.cpp needs boost::math). Neither is on the path-tracing hot path.
 * An in-memory raster, a minimal BMP reader and a reader for this repository's own OpenEXR files are provided;
 * every other file codec reports an error. */
#include <mitsuba/core/bitmap.h>
#include <mitsuba/render/testcase.h>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <sstream>
#include <vector>
#include <string>
#include <algorithm>
#include <zlib.h>

MTS_NAMESPACE_BEGIN

Bitmap::Bitmap(int width, int height, int bpp)
 : m_width(width), m_height(height), m_bpp(bpp), m_data(NULL) {
	m_gamma = (bpp == 96 || bpp == 128) ? 1.0f : -1.0f;
	m_size = (size_t) std::ceil(((double) m_width * m_height * m_bpp) / 8.0);
	m_data = static_cast<uint8_t *>(allocAligned(m_size));
}

/* The one codec the texture tests need: uncompressed 24-bit / 8-bit BMP without a palette, written from the BMP
 * format's definition (14-byte file header, 40-byte BITMAPINFOHEADER, rows padded to 4 bytes, bottom-up unless the
 * height is negative, BGR byte order). The result is what the reference's codec hands to its callers: top row
 * first, RGB order (src/libcore/bitmap.cpp:310-400). */
/* OpenEXR files as this repository's own writer produces them (scanline, HALF channels, no subsampling,
 * uncompressed / ZIPS / ZIP blocks), read from the published file layout: enough for the environment-map tests.
 * The result is what Bitmap::loadEXR hands on (src/libcore/bitmap.cpp:247-273): 128 bits per pixel, RGBA floats. */
static float halfToFloat(uint16_t h) {
	const uint32_t sign = (uint32_t) (h >> 15) << 31, exp = (h >> 10) & 31, man = h & 1023;
	uint32_t bits;
	if (exp == 0) {
		if (man == 0) bits = sign;
		else { int e = -1; uint32_t m = man; do { m <<= 1; ++e; } while (!(m & 1024)); bits = sign | ((uint32_t) (112 - e) << 23) | ((m & 1023) << 13); }
	} else if (exp == 31) bits = sign | 0x7F800000u | (man << 13);
	else bits = sign | ((exp + 112) << 23) | (man << 13);
	float f; memcpy(&f, &bits, 4); return f;
}

static void loadEXRSubset(Stream *stream, int &width, int &height, std::vector<float> &rgba) {
	std::vector<uint8_t> d(stream->getSize());
	stream->setPos(0); stream->read(d.data(), d.size());
	auto i32 = [&](size_t at) { int32_t v; memcpy(&v, &d[at], 4); return v; };
	if (d.size() < 8 || i32(0) != 20000630 || (i32(4) & 0xFF) != 2 || (i32(4) & ~0xFF))
		SLog(EError, "oracle build: not a plain scanline OpenEXR file");
	size_t pos = 8;
	std::vector<std::string> chans; int comp = -1, win[4] = { 0, 0, -1, -1 };
	while (d[pos] != 0) {
		std::string name((const char *) &d[pos]); pos += name.size() + 1;
		std::string type((const char *) &d[pos]); pos += type.size() + 1;
		const int size = i32(pos); pos += 4;
		if (name == "channels") {
			size_t p = pos;
			while (d[p] != 0) {
				std::string cn((const char *) &d[p]); p += cn.size() + 1;
				if (i32(p) != 1 || i32(p + 8) != 1 || i32(p + 12) != 1) SLog(EError, "oracle build: only HALF channels without subsampling");
				p += 16; chans.push_back(cn);
			}
		} else if (name == "compression") comp = d[pos];
		else if (name == "dataWindow") for (int k = 0; k < 4; ++k) win[k] = i32(pos + 4 * k);
		pos += size;
	}
	pos += 1;
	width = win[2] - win[0] + 1; height = win[3] - win[1] + 1;
	if (width <= 0 || height <= 0 || (comp != 0 && comp != 2 && comp != 3)) SLog(EError, "oracle build: unsupported OpenEXR file");
	const int lines = comp == 3 ? 16 : 1, nblk = (height + lines - 1) / lines, nc = (int) chans.size();
	rgba.assign((size_t) width * height * 4, 0.0f);
	for (size_t i = 0; i < (size_t) width * height; ++i) rgba[4 * i + 3] = 1.0f;
	for (int b = 0; b < nblk; ++b) {
		uint64_t off; memcpy(&off, &d[pos + 8 * (size_t) b], 8);
		const int y = i32(off), size = i32(off + 4), rows = std::min(lines, win[3] + 1 - y);
		const size_t rawSize = (size_t) rows * nc * width * 2;
		std::vector<uint8_t> raw(rawSize);
		if (comp != 0 && (size_t) size < rawSize) {
			std::vector<uint8_t> t(rawSize);
			uLongf got = rawSize;
			if (uncompress(t.data(), &got, &d[off + 8], size) != Z_OK || got != rawSize) SLog(EError, "oracle build: bad ZIP block");
			for (size_t k = 1; k < rawSize; ++k) t[k] = (uint8_t) (t[k - 1] + t[k] - 128);      /* delta predictor */
			const size_t half = (rawSize + 1) / 2;
			for (size_t k = 0; k < rawSize; ++k) raw[k] = (k & 1) ? t[half + k / 2] : t[k / 2];  /* de-interleave */
		} else memcpy(raw.data(), &d[off + 8], rawSize);
		for (int r = 0; r < rows; ++r)
			for (int c = 0; c < nc; ++c) {
				const int slot = chans[c] == "R" ? 0 : chans[c] == "G" ? 1 : chans[c] == "B" ? 2 : chans[c] == "A" ? 3 : -1;
				if (slot < 0) continue;
				for (int x = 0; x < width; ++x) {
					uint16_t h; memcpy(&h, &raw[(((size_t) r * nc + c) * width + x) * 2], 2);
					rgba[((size_t) (y - win[1] + r) * width + x) * 4 + slot] = halfToFloat(h);
				}
			}
	}
}

Bitmap::Bitmap(EFileFormat format, Stream *stream) : m_data(NULL) {
	if (format == EEXR) {
		std::vector<float> rgba;
		loadEXRSubset(stream, m_width, m_height, rgba);
		m_bpp = 128; m_gamma = 1.0f; m_size = rgba.size() * 4;
		m_data = static_cast<uint8_t *>(allocAligned(m_size));
		memcpy(m_data, rgba.data(), m_size);
		return;
	}
	if (format != EBMP)
		SLog(EError, "oracle build: only the BMP and OpenEXR codecs are available");
	uint8_t fh[14], ih[40];
	stream->read(fh, 14); stream->read(ih, 40);
	auto u32 = [](const uint8_t *p) { return (uint32_t) p[0] | ((uint32_t) p[1] << 8) | ((uint32_t) p[2] << 16) | ((uint32_t) p[3] << 24); };
	auto u16 = [](const uint8_t *p) { return (uint32_t) p[0] | ((uint32_t) p[1] << 8); };
	if (fh[0] != 'B' || fh[1] != 'M' || u32(ih) != 40 || u16(ih + 12) != 1 || u32(ih + 16) != 0 || u32(ih + 32) != 0)
		SLog(EError, "oracle build: unsupported BMP file");
	const int32_t h = (int32_t) u32(ih + 8);
	m_width = (int) u32(ih + 4); m_height = std::abs(h); m_bpp = (int) u16(ih + 14);
	if (m_bpp != 8 && m_bpp != 24)
		SLog(EError, "oracle build: only 8- and 24-bit BMP files");
	m_gamma = -1.0f;
	const int ch = m_bpp / 8, line = m_width * ch, padded = (line + 3) & ~3;
	m_size = (size_t) line * m_height;
	m_data = static_cast<uint8_t *>(allocAligned(m_size));
	stream->setPos(u32(fh + 10));
	std::vector<uint8_t> row(padded);
	for (int y = 0; y < m_height; ++y) {
		stream->read(row.data(), padded);
		uint8_t *dst = m_data + (size_t) (h > 0 ? m_height - 1 - y : y) * line;
		for (int x = 0; x < m_width; ++x)
			for (int c = 0; c < ch; ++c)
				dst[x * ch + c] = row[x * ch + (ch == 3 ? 2 - c : c)];
	}
}

Bitmap::~Bitmap() {
	if (m_data)
		freeAligned(m_data);
}

void Bitmap::clear() { memset(m_data, 0, m_size); }

Bitmap *Bitmap::clone() const {
	Bitmap *b = new Bitmap(m_width, m_height, m_bpp);
	memcpy(b->m_data, m_data, m_size);
	return b;
}

bool Bitmap::operator==(const Bitmap &o) const {
	return o.m_width == m_width && o.m_height == m_height && o.m_bpp == m_bpp
		&& memcmp(o.m_data, m_data, m_size) == 0;
}

void Bitmap::save(EFileFormat, Stream *, int) const {
	SLog(EError, "oracle build: image codecs are not available");
}

std::string Bitmap::toString() const {
	std::ostringstream oss;
	oss << "Bitmap[" << m_width << "x" << m_height << ", bpp=" << m_bpp << "]";
	return oss.str();
}

MTS_IMPLEMENT_CLASS(Bitmap, false, Object)

void TestSupervisor::analyze(const Scene *) {
	SLog(EError, "oracle build: TestSupervisor is not available");
}

MTS_NAMESPACE_END

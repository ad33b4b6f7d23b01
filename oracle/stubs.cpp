/* TEST INFRASTRUCTURE — link-time stand-ins for the two reference translation units that
 * cannot be compiled here (src/libcore/bitmap.cpp needs OpenEXR/libpng/libjpeg headers,
 * src/librender/testcase.cpp needs boost::math). Neither is on the path-tracing hot path.
 * An in-memory raster and a minimal BMP reader are provided; every other file codec reports an error. */
#include <mitsuba/core/bitmap.h>
#include <mitsuba/render/testcase.h>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <sstream>
#include <vector>

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
Bitmap::Bitmap(EFileFormat format, Stream *stream) : m_data(NULL) {
	if (format != EBMP)
		SLog(EError, "oracle build: only the BMP codec is available");
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

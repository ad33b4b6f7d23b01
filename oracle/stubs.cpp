/* TEST INFRASTRUCTURE — link-time stand-ins for the two reference translation units that
 * cannot be compiled here (src/libcore/bitmap.cpp needs OpenEXR/libpng/libjpeg headers,
 * src/librender/testcase.cpp needs boost::math). Neither is on the path-tracing hot path.
 * Only an in-memory raster is provided; every file codec reports an error.                */
#include <mitsuba/core/bitmap.h>
#include <mitsuba/render/testcase.h>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <sstream>

MTS_NAMESPACE_BEGIN

Bitmap::Bitmap(int width, int height, int bpp)
 : m_width(width), m_height(height), m_bpp(bpp), m_data(NULL) {
	m_gamma = (bpp == 96 || bpp == 128) ? 1.0f : -1.0f;
	m_size = (size_t) std::ceil(((double) m_width * m_height * m_bpp) / 8.0);
	m_data = static_cast<uint8_t *>(allocAligned(m_size));
}

Bitmap::Bitmap(EFileFormat, Stream *) : m_data(NULL) {
	SLog(EError, "oracle build: image codecs are not available");
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

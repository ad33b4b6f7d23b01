// Build shim (test infrastructure): the subset of boost::filesystem (v2 naming)
// the reference's hot-path sources use, expressed through std::filesystem.
#pragma once
#include <filesystem>
#include <string>
namespace boost { namespace filesystem {
class path : public std::filesystem::path {
public:
	typedef std::filesystem::path base;
	path() {}
	path(const base &p) : base(p) {}
	path(const std::string &s) : base(s) {}
	path(const char *s) : base(s) {}
	std::string file_string() const { return base::string(); }
	std::string leaf() const { return base::filename().string(); }
	path parent_path() const { return path(base::parent_path()); }
	path filename() const { return path(base::filename()); }
	std::string extension() const { return base::extension().string(); }
	path &replace_extension(const std::string &e = std::string()) { base::replace_extension(e); return *this; }
	path &operator/=(const path &o) { base::operator/=(static_cast<const base &>(o)); return *this; }
};
inline path operator/(const path &a, const path &b) { path r(a); r /= b; return r; }
inline bool exists(const path &p) { std::error_code ec; return std::filesystem::exists(p, ec); }
inline bool remove(const path &p) { std::error_code ec; return std::filesystem::remove(p, ec); }
inline unsigned long long file_size(const path &p) { return std::filesystem::file_size(p); }
inline path current_path() { return path(std::filesystem::current_path()); }
inline path complete(const path &p) { return path(std::filesystem::absolute(p)); }
inline std::string extension(const path &p) { return p.extension(); }
inline std::string basename(const path &p) { return p.stem().string(); }
} }

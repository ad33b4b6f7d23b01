// Build shim (test infrastructure): boost::filesystem stream types -> std.
#pragma once
#include <fstream>
#include <boost/filesystem.hpp>
namespace boost { namespace filesystem {
typedef std::ifstream ifstream;
typedef std::ofstream ofstream;
typedef std::fstream fstream;
} }

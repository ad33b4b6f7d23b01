// Build shim (test infrastructure): the two boost string helpers the reference uses.
#pragma once
#include <string>
#include <algorithm>
#include <cctype>
namespace boost {
inline std::string to_lower_copy(std::string s) {
	std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return (char) std::tolower(c); });
	return s;
}
inline bool ends_with(const std::string &s, const std::string &suffix) {
	return s.size() >= suffix.size() && s.compare(s.size() - suffix.size(), suffix.size(), suffix) == 0;
}
}

// Force-included when building oracle/_ref: the reference spells std::__cxx11::string (trajectory1d.cpp:196,320) and
// its header does not include <string>; libstdc++ only has std::__cxx11::basic_string under that name.  This makes
// the reference's spelling resolve to std::string.  Not reference code.
#pragma once
#include <string>
namespace std {
inline namespace __cxx11 {
using string = basic_string<char>;
}
}  // namespace std

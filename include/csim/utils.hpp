// utils.hpp — string / number helpers of the netlist front end.
// Same names and behaviour as the reference's include/utils.hpp (ltrim, toLower,
// parseSpiceNumber with the suffixes f p n u m k meg g t, isGroundName, clamp01).
#pragma once
#include <cctype>
#include <stdexcept>
#include <string>

inline std::string ltrim(const std::string& s) {
    std::size_t i = 0;
    while (i < s.size() && (s[i] == ' ' || s[i] == '\t' || s[i] == '\r' || s[i] == '\n')) ++i;
    return s.substr(i);
}

inline std::string toLower(const std::string& s) {
    std::string out(s);
    for (auto& ch : out) ch = static_cast<char>(std::tolower(static_cast<unsigned char>(ch)));
    return out;
}

namespace csim_detail {
inline double spiceSuffixScale(const std::string& suffix) {
    struct Entry {
        const char* tag;
        double scale;
    };
    static const Entry table[] = {{"f", 1e-15}, {"p", 1e-12}, {"n", 1e-9}, {"u", 1e-6}, {"m", 1e-3},
                                  {"k", 1e3},   {"meg", 1e6}, {"g", 1e9},  {"t", 1e12}};
    for (const Entry& e : table)
        if (suffix == e.tag) return e.scale;
    return 1.0;  // unknown suffix: ignored (reference: include/utils.hpp:42)
}
}  // namespace csim_detail

// "10k", "1u", "3e12", "3.3meg", ".25e-6" ...  (reference: include/utils.hpp:21-76)
// Numeric prefix by std::stod; whatever follows is looked up as a SPICE suffix.  A token
// with no leading number falls back to "text before the first letter"; no letters -> 0.
// May throw std::invalid_argument / std::out_of_range like the reference (callers catch).
inline double parseSpiceNumber(const std::string& token) {
    const std::string s = toLower(token);
    std::size_t used = 0;
    bool ok = true;
    double mantissa = 0.0;
    try {
        mantissa = std::stod(s, &used);
    } catch (...) {
        ok = false;
    }
    if (ok) {
        if (used == s.size()) return mantissa;
        return mantissa * csim_detail::spiceSuffixScale(s.substr(used));
    }
    std::size_t alpha = std::string::npos;
    for (std::size_t i = 0; i < s.size() && alpha == std::string::npos; ++i)
        if (std::isalpha(static_cast<unsigned char>(s[i]))) alpha = i;
    if (alpha == std::string::npos) return 0.0;
    mantissa = std::stod(s.substr(0, alpha));
    return mantissa * csim_detail::spiceSuffixScale(s.substr(alpha));
}

inline bool isGroundName(const std::string& n) {
    const std::string l = toLower(n);
    return l == "0" || l == "gnd";
}

inline double clamp01(double x) { return x < 0.0 ? 0.0 : (x > 1.0 ? 1.0 : x); }

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <string>

#include "variants.h"

namespace cusw {

std::vector<Variant> collect_variants() {
    std::vector<Variant> v;
    variants_g1(v); variants_g2(v); variants_g4(v); variants_g8(v); variants_g16(v); variants_g32(v);
    if (const char* e = getenv("CUDASW_VARIANTS")) {  // experiment hook: "8x19,4x38"
        std::vector<Variant> keep;
        std::string s(e);
        for (auto& x : v) {
            std::string tag = std::to_string(x.G) + "x" + std::to_string(x.R);
            size_t pos = 0;
            bool hit = false;
            while ((pos = s.find(tag, pos)) != std::string::npos) {
                bool l = pos == 0 || s[pos - 1] == ',';
                bool r = pos + tag.size() == s.size() || s[pos + tag.size()] == ',';
                if (l && r) { hit = true; break; }
                pos++;
            }
            if (hit) keep.push_back(x);
        }
        if (!keep.empty()) v = keep;
    }
    std::stable_sort(v.begin(), v.end(), [](const Variant& a, const Variant& b) {
        if (a.multi != b.multi) return b.multi;
        if (a.rows != b.rows) return a.rows < b.rows;
        const bool aw = a.R > 24, bw = b.R > 24;  // wide-R variants (low occupancy) only when nothing else fits as tightly
        if (aw != bw) return bw;
        return aw ? a.R < b.R : a.R > b.R;
    });
    // one variant per row count (the preferred one sorts first)
    std::vector<Variant> out;
    for (auto& x : v)
        if (out.empty() || out.back().rows != x.rows || out.back().multi != x.multi) out.push_back(x);
    return out;
}

}  // namespace cusw

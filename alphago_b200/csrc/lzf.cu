// LZF byte streams on the host: the chunk filter of the reference's dataset files (`compression="lzf"`,
// game_converter.py:73-74,81-82; h5py's registered HDF5 filter 32000, which wraps liblzf).  The format, restated from
// its definition: a stream of items, each starting with a control byte c --
//   c < 32        : a literal run of c + 1 bytes follows;
//   c >= 32       : a back reference: length = (c >> 5) (+ the next byte when that is 7) + 2, distance =
//                   ((c & 31) << 8 | next byte) + 1 bytes back from the output position (the copy may overlap itself).
// Declarations in include/alphago_b200.h.  Used by alphago_b200/hdf5_lite.py; nothing on the GPU path depends on it.
#include <stdint.h>
#include <string.h>

#include "../../include/alphago_b200.h"

namespace {

constexpr int kHashBits = 16;
constexpr unsigned kMaxDist = 1u << 13;   // 13 bits of distance
constexpr unsigned kMaxRef = (1u << 8) + (1u << 3);   // longest back reference: 7 + 255 + 2 = 264
constexpr unsigned kMaxLit = 1u << 5;     // longest literal run

inline uint32_t hash3(const uint8_t* p) {
    const uint32_t v = ((uint32_t)p[0] << 16) | ((uint32_t)p[1] << 8) | p[2];
    return ((v * 2654435761u) >> (32 - kHashBits)) & ((1u << kHashBits) - 1);
}

}  // namespace

extern "C" {

long long ag_lzf_decompress(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_cap) {
    size_t ip = 0, op = 0;
    while (ip < in_len) {
        unsigned ctrl = in[ip++];
        if (ctrl < 32) {
            const size_t n = ctrl + 1;
            if (ip + n > in_len || op + n > out_cap) return -1;
            memcpy(out + op, in + ip, n);
            ip += n; op += n;
        } else {
            size_t len = ctrl >> 5;
            if (len == 7) {
                if (ip >= in_len) return -1;
                len += in[ip++];
            }
            if (ip >= in_len) return -1;
            const size_t dist = (((size_t)(ctrl & 31)) << 8 | in[ip++]) + 1;
            len += 2;
            if (dist > op || op + len > out_cap) return -1;
            const uint8_t* ref = out + op - dist;
            for (size_t i = 0; i < len; i++) out[op + i] = ref[i];   // byte by byte: the run may overlap its source
            op += len;
        }
    }
    return (long long)op;
}

long long ag_lzf_compress(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_cap) {
    if (in_len == 0) return 0;
    static thread_local uint32_t table[1u << kHashBits];   // position + 1 of the last occurrence of a 3-byte hash
    memset(table, 0, sizeof(table));
    size_t ip = 0, op = 0, lit_start = 0;
    auto flush_literals = [&](size_t end) -> bool {          // bytes [lit_start, end) as literal runs
        while (lit_start < end) {
            const size_t n = end - lit_start < kMaxLit ? end - lit_start : kMaxLit;
            if (op + 1 + n > out_cap) return false;
            out[op++] = (uint8_t)(n - 1);
            memcpy(out + op, in + lit_start, n);
            op += n; lit_start += n;
        }
        return true;
    };
    while (ip + 2 < in_len) {
        const uint32_t h = hash3(in + ip);
        const size_t cand = table[h];
        table[h] = (uint32_t)(ip + 1);
        if (cand && ip - (cand - 1) <= kMaxDist && in[cand - 1] == in[ip] && in[cand] == in[ip + 1] && in[cand + 1] == in[ip + 2]) {
            const size_t ref = cand - 1;
            size_t len = 3;
            const size_t max_len = in_len - ip < kMaxRef ? in_len - ip : kMaxRef;
            while (len < max_len && in[ref + len] == in[ip + len]) len++;
            if (!flush_literals(ip)) return 0;
            const size_t dist = ip - ref - 1, l = len - 2;
            if (op + 3 > out_cap) return 0;
            if (l < 7) out[op++] = (uint8_t)((l << 5) | (dist >> 8));
            else { out[op++] = (uint8_t)((7u << 5) | (dist >> 8)); out[op++] = (uint8_t)(l - 7); }
            out[op++] = (uint8_t)(dist & 255);
            // index the match's second position and its last two: enough for the next match to find its way back, and long
            // runs (these planes are mostly zeros) cost three table writes instead of hundreds
            const size_t end = ip + len;
            if (ip + 3 < in_len) table[hash3(in + ip + 1)] = (uint32_t)(ip + 2);
            for (size_t q = end - 2; q < end; q++)
                if (q > ip + 1 && q + 2 < in_len) table[hash3(in + q)] = (uint32_t)(q + 1);
            ip = end;
            lit_start = ip;
        } else {
            ip++;
        }
    }
    if (!flush_literals(in_len)) return 0;
    return (long long)op;
}

}  // extern "C"

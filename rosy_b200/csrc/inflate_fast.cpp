// inflate_fast.cpp - DEFLATE (RFC 1951) / zlib (RFC 1950) decoder for the PNG loader (nvtt::Surface::load,
// /root/reference/src/Packager/Packager.cpp:222,269).
//
// The file pipeline is decode-bound: one core spent 123 of the 238 ms of a 2048^2 PNG inside
// zlib's inflate.  This decoder is written for that case - the whole compressed stream and the whole output buffer in
// memory, output size known in advance - which removes zlib's streaming state machine: a 64-bit bit buffer refilled
// with one unaligned load, an 11-bit litlen table whose entries carry literal / length base / extra-bit count ready to
// use, up to three literals per refill, matches copied eight bytes at a time.  Written from the RFCs; the table layout
// follows the well-known "decode table entry = everything the hot loop needs" design of modern inflate implementations.
// Anything unusual (preset dictionary, invalid code, overrun) returns false and the caller falls back to zlib, which
// then produces the error or the data, so behaviour on malformed files is zlib's.
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#ifdef INFLATE_FAST_DEBUG
#include <stdio.h>
#include <stdlib.h>
#endif

namespace rosy_tex {
namespace {

constexpr int LITLEN_BITS = 11, OFFSET_BITS = 8, MAX_CODE_LEN = 15;
constexpr int LITLEN_SYMS = 288, OFFSET_SYMS = 32, PRECODE_SYMS = 19;
// entry: low 8 bits = bits to consume now; bits 8..15 = extra bits (lengths, offsets) or subtable width;
// bits 16..30 = literal / base value / subtable start; flags below
// (offset entries need 15 bits of base value: there the sub-table flag sits in bit 31 and the value in bits 16..30)
constexpr uint32_t F_LITERAL = 0x80000000u, F_SUBTABLE = 0x40000000u, F_EOB = 0x20000000u, F_VALUE_MASK = 0x1FFF0000u;
constexpr uint32_t O_SUBTABLE = 0x80000000u, O_VALUE_MASK = 0x7FFF0000u;
constexpr int LITLEN_ENTRIES = (1 << LITLEN_BITS) + LITLEN_SYMS * (1 << (MAX_CODE_LEN - LITLEN_BITS));
constexpr int OFFSET_ENTRIES = (1 << OFFSET_BITS) + OFFSET_SYMS * (1 << (MAX_CODE_LEN - OFFSET_BITS));

const uint16_t kLenBase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
const uint8_t kLenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
const uint16_t kOffBase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
const uint8_t kOffExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
const uint8_t kPrecodeOrder[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};

inline uint32_t reverse_bits(uint32_t code, int len) {
    uint32_t r = 0;
    for (int i = 0; i < len; i++) r |= ((code >> i) & 1u) << (len - 1 - i);
    return r;
}

// what a symbol's table entry holds besides the bit count
constexpr uint32_t INVALID = 0xFFFFFFFFu;  // symbols 286, 287 / offsets 30, 31: part of the fixed code, never legal in a stream
inline uint32_t litlen_payload(int sym) {
    if (sym < 256) return F_LITERAL | ((uint32_t)sym << 16);
    if (sym == 256) return F_EOB;
    const int k = sym - 257;
    if (k >= 29) return INVALID;
    return ((uint32_t)kLenBase[k] << 16) | ((uint32_t)kLenExtra[k] << 8);
}
inline uint32_t offset_payload(int sym) {
    if (sym >= 30) return INVALID;
    return ((uint32_t)kOffBase[sym] << 16) | ((uint32_t)kOffExtra[sym] << 8);
}

// Canonical Huffman code -> decode table (primary table of `tbits` index bits + subtables of max_len - tbits bits).
// Accepts complete codes, and (as RFC 1951 allows for distances) a code with a single symbol of length 1.
template <bool LITLEN>
bool build_table(const uint8_t* lens, int nsyms, int tbits, uint32_t* table, int table_cap) {
    int count[MAX_CODE_LEN + 1] = {0};
    for (int s = 0; s < nsyms; s++) count[lens[s]]++;
    int max_len = MAX_CODE_LEN;
    while (max_len > 0 && count[max_len] == 0) max_len--;
    if (max_len == 0) {
        if (LITLEN) return false;  // a litlen code needs at least the end-of-block symbol
        for (int i = 0; i < (1 << tbits); i++) table[i] = 0;  // no distance codes: any use of one is an error (bits = 0)
        return true;
    }
    long left = 1;  // Kraft
    for (int len = 1; len <= max_len; len++) {
        left = (left << 1) - count[len];
        if (left < 0) return false;  // over-subscribed
    }
    const bool single = !LITLEN && max_len == 1 && count[1] == 1;
    if (left != 0 && !single) return false;  // incomplete codes other than the one-distance case: let zlib judge
    uint32_t next_code[MAX_CODE_LEN + 2];
    uint32_t code = 0;
    for (int len = 1; len <= max_len; len++) {
        next_code[len] = code;
        code = (code + (uint32_t)count[len]) << 1;
    }
    const int sub_bits = max_len > tbits ? max_len - tbits : 0;
    int next_sub = 1 << tbits;
    for (int i = 0; i < (1 << tbits); i++) table[i] = 0;
    for (int s = 0; s < nsyms; s++) {
        const int len = lens[s];
        if (!len) continue;
        const uint32_t rev = reverse_bits(next_code[len]++, len);
        const uint32_t payload = LITLEN ? litlen_payload(s) : offset_payload(s);
        if (payload == INVALID) continue;  // its code points keep the error entry (bits = 0)
        if (len <= tbits) {
            for (uint32_t i = rev; i < (1u << tbits); i += 1u << len) table[i] = payload | (uint32_t)len;
        } else {
            const uint32_t prefix = rev & ((1u << tbits) - 1u);
            constexpr uint32_t SUB = LITLEN ? F_SUBTABLE : O_SUBTABLE, VMASK = LITLEN ? F_VALUE_MASK : O_VALUE_MASK;
            if (!(table[prefix] & SUB)) {
                if (next_sub + (1 << sub_bits) > table_cap) return false;
                table[prefix] = SUB | ((uint32_t)next_sub << 16) | ((uint32_t)sub_bits << 8) | (uint32_t)tbits;
                for (int i = 0; i < (1 << sub_bits); i++) table[next_sub + i] = 0;
                next_sub += 1 << sub_bits;
            }
            const uint32_t base = (table[prefix] & VMASK) >> 16;
            for (uint32_t i = rev >> tbits; i < (1u << sub_bits); i += 1u << (len - tbits))
                table[base + i] = payload | (uint32_t)(len - tbits);
        }
    }
    return true;  // the unused half of a one-symbol distance code keeps the error entry (bits = 0)
}

struct Tables {
    uint32_t litlen[LITLEN_ENTRIES];
    uint32_t offset[OFFSET_ENTRIES];
};

inline uint64_t load64(const uint8_t* p) {
    uint64_t v;
    memcpy(&v, p, 8);
    return v;  // little endian hosts only (x86-64, aarch64)
}

}  // namespace

// in: `in_len` bytes of a zlib stream (raw = false) or a raw deflate stream (raw = true), FOLLOWED BY at least 32
// readable padding bytes.  out: capacity out_len + 16 (wide copies may write up to 7 bytes past a match, a run of
// literals up to 3 past the end before the overrun is seen).  Succeeds only if
// the stream is well formed and inflates to exactly out_len bytes; the zlib trailer (Adler-32) is not checked (stb_image,
// which the reference links, does not check it either).
bool inflate_fast(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_len, bool raw) {
    static thread_local Tables T;
    const uint8_t* const in_end = in + in_len;
    uint8_t* const out_begin = out;
    uint8_t* const out_end = out + out_len;
    if (!raw) {
        if (in_len < 6) return false;
        const unsigned cmf = in[0], flg = in[1];
        if ((cmf & 15u) != 8u || (cmf >> 4) > 7u || ((cmf << 8) | flg) % 31u != 0u || (flg & 32u)) return false;
        in += 2;
    }
    uint64_t bitbuf = 0;
    unsigned bitcnt = 0;
    // refill to at least 56 valid bits; reads 8 bytes at `in` (the caller's padding makes that safe at the very end)
#define REFILL()                                  \
    do {                                          \
        bitbuf |= load64(in) << bitcnt;           \
        in += (63u - bitcnt) >> 3;                \
        bitcnt |= 56u;                            \
    } while (0)
#ifdef INFLATE_FAST_DEBUG
#define CONSUME(n) do { if ((unsigned)(n) > bitcnt) { fprintf(stderr, "underflow line %d n %u bitcnt %u\n", __LINE__, (unsigned)(n), bitcnt); abort(); } bitbuf >>= (n); bitcnt -= (n); } while (0)
#else
#define CONSUME(n) (bitbuf >>= (n), bitcnt -= (n))
#endif
    // bits of the padding have been consumed: the stream is truncated (the 32 padding bytes cover the loads until this
    // is tested again: without an overrun `in` is at most 7 bytes past the end before a refill, see REFILL)
#define OVERRUN() (in > in_end && (size_t)(in - in_end) * 8u > bitcnt)
    bool last = false;
    while (!last) {
        if (OVERRUN()) return false;
        REFILL();
        last = (bitbuf & 1u) != 0;
        const unsigned type = (unsigned)(bitbuf >> 1) & 3u;
        CONSUME(3);
        if (type == 0) {  // stored: back to a byte boundary, LEN / NLEN, raw bytes
            CONSUME(bitcnt & 7u);
            in -= bitcnt >> 3;  // un-read the whole bytes still in the buffer
            bitbuf = 0;
            bitcnt = 0;
            if (in + 4 > in_end) return false;
            const unsigned len = in[0] | (in[1] << 8), nlen = in[2] | (in[3] << 8);
            if ((len ^ nlen) != 0xFFFFu) return false;
            in += 4;
            if ((size_t)(in_end - in) < len || (size_t)(out_end - out) < len) return false;
            memcpy(out, in, len);
            in += len;
            out += len;
            continue;
        }
        if (type == 3) return false;
        uint8_t lens[LITLEN_SYMS + OFFSET_SYMS];
        if (type == 1) {  // fixed code
            for (int i = 0; i < 144; i++) lens[i] = 8;
            for (int i = 144; i < 256; i++) lens[i] = 9;
            for (int i = 256; i < 280; i++) lens[i] = 7;
            for (int i = 280; i < 288; i++) lens[i] = 8;
            uint8_t olens[OFFSET_SYMS];
            for (int i = 0; i < 32; i++) olens[i] = 5;
            if (!build_table<true>(lens, LITLEN_SYMS, LITLEN_BITS, T.litlen, LITLEN_ENTRIES)) return false;
            if (!build_table<false>(olens, OFFSET_SYMS, OFFSET_BITS, T.offset, OFFSET_ENTRIES)) return false;
        } else {  // dynamic code
            const unsigned hlit = ((unsigned)bitbuf & 31u) + 257u, hdist = ((unsigned)(bitbuf >> 5) & 31u) + 1u,
                           hclen = ((unsigned)(bitbuf >> 10) & 15u) + 4u;
            CONSUME(14);
            if (hlit > 286u || hdist > 30u) return false;
            uint8_t plens[PRECODE_SYMS] = {0};
            for (unsigned i = 0; i < hclen; i++) {
                if (bitcnt < 3) {
                    if (OVERRUN()) return false;
                    REFILL();
                }
                plens[kPrecodeOrder[i]] = (uint8_t)(bitbuf & 7u);
                CONSUME(3);
            }
            // the code-length code: at most 7 bits, a plain 128-entry table (symbol << 8 | length)
            uint16_t ptab[128];
            {
                int count[8] = {0};
                for (int s = 0; s < PRECODE_SYMS; s++) count[plens[s]]++;
                long left = 1;
                for (int len = 1; len <= 7; len++) {
                    left = (left << 1) - count[len];
                    if (left < 0) return false;
                }
                if (left != 0) return false;
                uint32_t next_code[9], code = 0;
                for (int len = 1; len <= 7; len++) {
                    next_code[len] = code;
                    code = (code + (uint32_t)count[len]) << 1;
                }
                for (int s = 0; s < PRECODE_SYMS; s++) {
                    const int len = plens[s];
                    if (!len) continue;
                    const uint32_t rev = reverse_bits(next_code[len]++, len);
                    for (uint32_t i = rev; i < 128u; i += 1u << len) ptab[i] = (uint16_t)((s << 8) | len);
                }
            }
            unsigned n = 0;
            while (n < hlit + hdist) {
                if (OVERRUN()) return false;
                REFILL();
                const uint16_t e = ptab[bitbuf & 127u];
                CONSUME(e & 255u);
                const unsigned sym = e >> 8;
                if (sym < 16) {
                    lens[n++] = (uint8_t)sym;
                    continue;
                }
                unsigned rep, val = 0;
                if (sym == 16) {
                    if (n == 0) return false;
                    val = lens[n - 1];
                    rep = 3u + ((unsigned)bitbuf & 3u);
                    CONSUME(2);
                } else if (sym == 17) {
                    rep = 3u + ((unsigned)bitbuf & 7u);
                    CONSUME(3);
                } else {
                    rep = 11u + ((unsigned)bitbuf & 127u);
                    CONSUME(7);
                }
                if (n + rep > hlit + hdist) return false;
                memset(lens + n, (int)val, rep);
                n += rep;
            }
            if (lens[256] == 0) return false;  // no end-of-block code
            uint8_t ll[LITLEN_SYMS] = {0}, ol[OFFSET_SYMS] = {0};
            memcpy(ll, lens, hlit);
            memcpy(ol, lens + hlit, hdist);
            if (!build_table<true>(ll, LITLEN_SYMS, LITLEN_BITS, T.litlen, LITLEN_ENTRIES)) return false;
            if (!build_table<false>(ol, OFFSET_SYMS, OFFSET_BITS, T.offset, OFFSET_ENTRIES)) return false;
        }
        // ---- the block's symbols ----
        for (;;) {
            if (OVERRUN() || out > out_end) return false;
            REFILL();
            uint32_t e = T.litlen[bitbuf & ((1u << LITLEN_BITS) - 1u)];
            // up to three literals per refill (3 x 15 bits <= 56 when none needs a subtable; a subtable entry ends the run)
            if (e & F_LITERAL) {
                CONSUME(e & 255u);
                *out++ = (uint8_t)(e >> 16);
                e = T.litlen[bitbuf & ((1u << LITLEN_BITS) - 1u)];
                if (e & F_LITERAL) {
                    CONSUME(e & 255u);
                    *out++ = (uint8_t)(e >> 16);
                    e = T.litlen[bitbuf & ((1u << LITLEN_BITS) - 1u)];
                    if (e & F_LITERAL) {
                        CONSUME(e & 255u);
                        *out++ = (uint8_t)(e >> 16);
                        continue;
                    }
                }
                // two literals from the primary table took at most 22 bits: at least 34 are left for the length below
            }
            if (e & F_SUBTABLE) {
                CONSUME(e & 255u);
                e = T.litlen[((e & F_VALUE_MASK) >> 16) + (bitbuf & ((1u << ((e >> 8) & 15u)) - 1u))];
                if (e & F_LITERAL) {
                    CONSUME(e & 255u);
                    *out++ = (uint8_t)(e >> 16);
                    continue;
                }
            }
            if ((e & 255u) == 0 || out > out_end) return false;  // invalid code, or the literals above ran past the end
            CONSUME(e & 255u);
            if (e & F_EOB) break;
            const unsigned lextra = (e >> 8) & 31u;
            unsigned length = ((e & F_VALUE_MASK) >> 16) + ((unsigned)bitbuf & ((1u << lextra) - 1u));
            CONSUME(lextra);
            if (bitcnt < 32) REFILL();
            uint32_t o = T.offset[bitbuf & ((1u << OFFSET_BITS) - 1u)];
            if (o & O_SUBTABLE) {
                CONSUME(o & 255u);
                o = T.offset[((o & O_VALUE_MASK) >> 16) + (bitbuf & ((1u << ((o >> 8) & 15u)) - 1u))];
            }
            if ((o & 255u) == 0) return false;
            CONSUME(o & 255u);
            const unsigned oextra = (o >> 8) & 31u;
            const size_t dist = ((o & O_VALUE_MASK) >> 16) + ((unsigned)bitbuf & ((1u << oextra) - 1u));
            CONSUME(oextra);
            if (dist > (size_t)(out - out_begin) || length > (size_t)(out_end - out)) return false;
            const uint8_t* src = out - dist;
            uint8_t* dst = out;
            out += length;
            if (dist >= 8) {
                // eight bytes at a time; may write up to 7 bytes past the match (the caller's slack)
                do {
                    memcpy(dst, src, 8);
                    dst += 8;
                    src += 8;
                } while (dst < out);
            } else if (dist == 1) {
                memset(dst, src[0], length);
            } else {
                do {
                    *dst++ = *src++;
                } while (dst < out);
            }
        }
    }
    const bool ok = out == out_end && !OVERRUN();
#undef REFILL
#undef CONSUME
#undef OVERRUN
    return ok;
}

}  // namespace rosy_tex

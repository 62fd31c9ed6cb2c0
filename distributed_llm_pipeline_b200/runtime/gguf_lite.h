// Minimal GGUF v2/v3 reader shared by the native runtime binaries (dlp-worker, dlp-run): mmap, KV scalars / strings / array
// lengths, tensor directory, and scalar row decoders for the tensor types the runtime handles (needs -mf16c).  Format: llama.cpp/ggml/include/gguf.h:1-61, reader llama.cpp/ggml/src/gguf.cpp.
#pragma once
#include <fcntl.h>
#include <immintrin.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstdint>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace dlp_gguf {

enum GType : uint32_t { T_F32 = 0, T_F16 = 1, T_Q4_0 = 2, T_Q4_1 = 3, T_Q5_0 = 6, T_Q5_1 = 7, T_Q8_0 = 8, T_Q2_K = 10, T_Q3_K = 11, T_Q4_K = 12, T_Q5_K = 13,
                        T_Q6_K = 14, T_IQ4_NL = 20, T_IQ4_XS = 23, T_BF16 = 30 };

// ---- GGUF -----------------------------------------------------------------------------------------------------------
struct Tensor {
    std::vector<uint64_t> ne;  // innermost first
    uint32_t type = 0;
    const uint8_t* data = nullptr;
    uint64_t rows() const { uint64_t r = 1; for (size_t i = 1; i < ne.size(); ++i) r *= ne[i]; return r; }
    uint64_t cols() const { return ne.empty() ? 0 : ne[0]; }
};

inline size_t row_bytes(uint32_t type, uint64_t k);
// rows [r0, r1) of a 2-D tensor as a tensor of its own (phi3 keeps Q|K|V and gate|up in one matrix each)
inline Tensor slice_rows(const Tensor& t, uint64_t r0, uint64_t r1) {
    Tensor v;
    v.ne = {t.cols(), r1 - r0};
    v.type = t.type;
    v.data = t.data + r0 * row_bytes(t.type, t.cols());
    return v;
}

inline size_t row_bytes(uint32_t type, uint64_t k) {
    switch (type) {
        case T_F32: return k * 4;
        case T_F16: case T_BF16: return k * 2;
        case T_Q8_0: return k / 32 * 34;
        case T_Q4_0: case T_IQ4_NL: return k / 32 * 18;
        case T_Q4_1: return k / 32 * 20;
        case T_Q5_0: return k / 32 * 22;
        case T_Q5_1: return k / 32 * 24;
        case T_Q2_K: return k / 256 * 84;
        case T_Q3_K: return k / 256 * 110;
        case T_IQ4_XS: return k / 256 * 136;
        case T_Q4_K: return k / 256 * 144;
        case T_Q5_K: return k / 256 * 176;
        case T_Q6_K: return k / 256 * 210;
        default: throw std::runtime_error("unsupported tensor type " + std::to_string(type));
    }
}

struct Gguf {
    const uint8_t* base = nullptr;
    size_t size = 0;
    std::map<std::string, double> num;
    std::map<std::string, std::string> str;
    std::map<std::string, uint64_t> arr_len;
    std::map<std::string, Tensor> tensors;

    struct Cur {
        const uint8_t* p; const uint8_t* end;
        template <typename T> T get() { if (p + sizeof(T) > end) throw std::runtime_error("gguf: truncated"); T v; memcpy(&v, p, sizeof(T)); p += sizeof(T); return v; }
        std::string s() { uint64_t n = get<uint64_t>(); if (p + n > end) throw std::runtime_error("gguf: truncated string"); std::string r((const char*)p, n); p += n; return r; }
    };
    static size_t scalar_size(uint32_t t) { static const size_t sz[13] = {1, 1, 2, 2, 4, 4, 4, 1, 0, 0, 8, 8, 8}; return t < 13 ? sz[t] : 0; }

    void skip_or_read(Cur& c, uint32_t t, const std::string& key) {
        if (t == 8) { str[key] = c.s(); return; }
        if (t == 9) {
            uint32_t et = c.get<uint32_t>(); uint64_t n = c.get<uint64_t>();
            arr_len[key] = n;
            if (et == 8) { for (uint64_t i = 0; i < n; ++i) c.s(); }
            else { size_t s = scalar_size(et); if (!s) throw std::runtime_error("gguf: nested arrays unsupported"); if (c.p + s * n > c.end) throw std::runtime_error("gguf: truncated array"); c.p += s * n; }
            return;
        }
        double v = 0;
        switch (t) {
            case 0: v = c.get<uint8_t>(); break;  case 1: v = c.get<int8_t>(); break;
            case 2: v = c.get<uint16_t>(); break; case 3: v = c.get<int16_t>(); break;
            case 4: v = c.get<uint32_t>(); break; case 5: v = c.get<int32_t>(); break;
            case 6: v = c.get<float>(); break;    case 7: v = c.get<uint8_t>(); break;
            case 10: v = (double)c.get<uint64_t>(); break; case 11: v = (double)c.get<int64_t>(); break;
            case 12: v = c.get<double>(); break;
            default: throw std::runtime_error("gguf: bad value type");
        }
        num[key] = v;
    }

    void open(const std::string& path) {
        int fd = ::open(path.c_str(), O_RDONLY);
        if (fd < 0) throw std::runtime_error("cannot open " + path);
        struct stat st{};
        fstat(fd, &st);
        size = (size_t)st.st_size;
        base = (const uint8_t*)mmap(nullptr, size, PROT_READ, MAP_SHARED, fd, 0);
        ::close(fd);
        if (base == MAP_FAILED) throw std::runtime_error("mmap failed");
        Cur c{base, base + size};
        if (c.get<uint32_t>() != 0x46554747u) throw std::runtime_error("not a GGUF file");
        uint32_t ver = c.get<uint32_t>();
        if (ver != 2 && ver != 3) throw std::runtime_error("unsupported GGUF version");
        uint64_t nt = c.get<uint64_t>(), nkv = c.get<uint64_t>();
        for (uint64_t i = 0; i < nkv; ++i) { std::string k = c.s(); uint32_t t = c.get<uint32_t>(); skip_or_read(c, t, k); }
        struct Info { std::string name; Tensor t; uint64_t off; };
        std::vector<Info> infos;
        for (uint64_t i = 0; i < nt; ++i) {
            Info in; in.name = c.s();
            uint32_t nd = c.get<uint32_t>();
            for (uint32_t d = 0; d < nd; ++d) in.t.ne.push_back(c.get<uint64_t>());
            in.t.type = c.get<uint32_t>(); in.off = c.get<uint64_t>();
            infos.push_back(in);
        }
        uint64_t align = num.count("general.alignment") ? (uint64_t)num["general.alignment"] : 32;
        uint64_t data_off = ((uint64_t)(c.p - base) + align - 1) / align * align;
        for (auto& in : infos) {
            in.t.data = base + data_off + in.off;
            tensors[in.name] = in.t;
        }
    }
    const Tensor& get(const std::string& n) const { auto it = tensors.find(n); if (it == tensors.end()) throw std::runtime_error("missing tensor " + n); return it->second; }
    bool has(const std::string& n) const { return tensors.count(n) > 0; }
    double n(const std::string& k, double d) const { auto it = num.find(k); return it == num.end() ? d : it->second; }
};

// ---- row decoding ----------------------------------------------------------------------------------------------------
inline float f16(uint16_t h) { return _cvtsh_ss(h); }

inline void scale_min_k4(int j, const uint8_t* q, int& d, int& m) {
    if (j < 4) { d = q[j] & 63; m = q[j + 4] & 63; }
    else { d = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4); m = (q[j + 4] >> 4) | ((q[j] >> 6) << 4); }
}

inline void decode_row(uint32_t type, const uint8_t* src, uint64_t k, float* out) {
    switch (type) {
        case T_F32: memcpy(out, src, k * 4); break;
        case T_F16: { const uint16_t* h = (const uint16_t*)src; for (uint64_t i = 0; i < k; ++i) out[i] = f16(h[i]); break; }
        case T_BF16: { const uint16_t* h = (const uint16_t*)src; for (uint64_t i = 0; i < k; ++i) { uint32_t u = (uint32_t)h[i] << 16; memcpy(out + i, &u, 4); } break; }
        case T_Q8_0:
            for (uint64_t b = 0; b < k / 32; ++b) {
                const uint8_t* p = src + b * 34;
                uint16_t dh; memcpy(&dh, p, 2);
                const float d = f16(dh);
                const int8_t* q = (const int8_t*)(p + 2);
                for (int i = 0; i < 32; ++i) out[b * 32 + i] = d * q[i];
            }
            break;
        case T_Q4_K:
            for (uint64_t b = 0; b < k / 256; ++b) {
                const uint8_t* p = src + b * 144;
                uint16_t dh, mh; memcpy(&dh, p, 2); memcpy(&mh, p + 2, 2);
                const float d = f16(dh), dm = f16(mh);
                const uint8_t* sc = p + 4; const uint8_t* q = p + 16;
                float* y = out + b * 256;
                for (int j = 0; j < 4; ++j) {
                    int s1, m1, s2, m2;
                    scale_min_k4(2 * j, sc, s1, m1); scale_min_k4(2 * j + 1, sc, s2, m2);
                    for (int l = 0; l < 32; ++l) { y[l] = d * s1 * (q[l] & 0xF) - dm * m1; y[l + 32] = d * s2 * (q[l] >> 4) - dm * m2; }
                    y += 64; q += 32;
                }
            }
            break;
        case T_Q5_K:
            for (uint64_t b = 0; b < k / 256; ++b) {
                const uint8_t* p = src + b * 176;
                uint16_t dh, mh; memcpy(&dh, p, 2); memcpy(&mh, p + 2, 2);
                const float d = f16(dh), dm = f16(mh);
                const uint8_t* sc = p + 4; const uint8_t* qh = p + 16; const uint8_t* q = p + 48;
                float* y = out + b * 256;
                for (int j = 0; j < 4; ++j) {
                    int s1, m1, s2, m2;
                    scale_min_k4(2 * j, sc, s1, m1); scale_min_k4(2 * j + 1, sc, s2, m2);
                    for (int l = 0; l < 32; ++l) {
                        y[l] = d * s1 * ((q[l] & 0xF) + (((qh[l] >> (2 * j)) & 1) << 4)) - dm * m1;
                        y[l + 32] = d * s2 * ((q[l] >> 4) + (((qh[l] >> (2 * j + 1)) & 1) << 4)) - dm * m2;
                    }
                    y += 64; q += 32;
                }
            }
            break;
        case T_Q6_K:
            for (uint64_t b = 0; b < k / 256; ++b) {
                const uint8_t* p = src + b * 210;
                const uint8_t* ql = p; const uint8_t* qh = p + 128; const int8_t* sc = (const int8_t*)(p + 192);
                uint16_t dh; memcpy(&dh, p + 208, 2);
                const float d = f16(dh);
                float* y = out + b * 256;
                for (int half = 0; half < 2; ++half) {
                    for (int l = 0; l < 32; ++l) {
                        const int is = l / 16;
                        const int q1 = (int)((ql[l] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                        const int q2 = (int)((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                        const int q3 = (int)((ql[l] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                        const int q4 = (int)((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
                        y[l] = d * sc[is] * q1; y[l + 32] = d * sc[is + 2] * q2; y[l + 64] = d * sc[is + 4] * q3; y[l + 96] = d * sc[is + 6] * q4;
                    }
                    y += 128; ql += 64; qh += 32; sc += 8;
                }
            }
            break;
        // legacy 4- and 5-bit blocks of 32: fp16 scale (and minimum for the _1 kinds), the fifth bits of Q5 in one little-endian word,
        // nibbles with element j in the low half of byte j and element j + 16 in the high half
        case T_Q4_0: case T_Q4_1: case T_Q5_0: case T_Q5_1: case T_IQ4_NL: {
            static const float iq4[16] = {-127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113};
            const bool has_min = type == T_Q4_1 || type == T_Q5_1, five = type == T_Q5_0 || type == T_Q5_1;
            const size_t bs = row_bytes(type, 32);
            for (uint64_t b = 0; b < k / 32; ++b) {
                const uint8_t* p = src + b * bs;
                uint16_t dh, mh = 0; memcpy(&dh, p, 2); p += 2;
                if (has_min) { memcpy(&mh, p, 2); p += 2; }
                uint32_t hb = 0;
                if (five) { memcpy(&hb, p, 4); p += 4; }
                const float d = f16(dh), m = has_min ? f16(mh) : 0.0f;
                float* y = out + b * 32;
                for (int i = 0; i < 16; ++i) {
                    int lo = p[i] & 0xF, hi = p[i] >> 4;
                    if (five) { lo |= ((hb >> i) & 1) << 4; hi |= ((hb >> (i + 16)) & 1) << 4; }
                    if (type == T_IQ4_NL) { y[i] = d * iq4[lo]; y[i + 16] = d * iq4[hi]; }
                    else if (has_min) { y[i] = d * lo + m; y[i + 16] = d * hi + m; }
                    else { const int off = five ? 16 : 8; y[i] = d * (lo - off); y[i + 16] = d * (hi - off); }
                }
            }
            break;
        }
        case T_IQ4_XS:      // 256 values: fp16 scale, eight 6-bit sub-block scales (two high bits in one word, low nibbles in 4 bytes), non-linear nibbles
            for (uint64_t b = 0; b < k / 256; ++b) {
                static const float iq4[16] = {-127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113};
                const uint8_t* p = src + b * 136;
                uint16_t dh, sh; memcpy(&dh, p, 2); memcpy(&sh, p + 2, 2);
                const float d = f16(dh);
                const uint8_t* sl = p + 4; const uint8_t* q = p + 8;
                float* y = out + b * 256;
                for (int ib = 0; ib < 8; ++ib) {
                    const int ls = (((sl[ib / 2] >> (4 * (ib % 2))) & 0xF) | (((sh >> (2 * ib)) & 3) << 4)) - 32;
                    const float dl = d * ls;
                    for (int i = 0; i < 16; ++i) { y[i] = dl * iq4[q[i] & 0xF]; y[i + 16] = dl * iq4[q[i] >> 4]; }
                    y += 32; q += 16;
                }
            }
            break;
        case T_Q2_K:        // 16 sub-blocks of 16: 4-bit scale | 4-bit minimum per sub-block, 2-bit values in four bit planes per 32 bytes
            for (uint64_t b = 0; b < k / 256; ++b) {
                const uint8_t* p = src + b * 84;
                const uint8_t* sc = p; const uint8_t* q = p + 16;
                uint16_t dh, mh; memcpy(&dh, p + 80, 2); memcpy(&mh, p + 82, 2);
                const float d = f16(dh), dm = f16(mh);
                float* y = out + b * 256;
                int is = 0;
                for (int half = 0; half < 2; ++half, q += 32)
                    for (int sh = 0; sh < 8; sh += 2)
                        for (int part = 0; part < 2; ++part, ++is) {
                            const float dl = d * (sc[is] & 0xF), ml = dm * (sc[is] >> 4);
                            for (int l = 0; l < 16; ++l) *y++ = dl * ((q[part * 16 + l] >> sh) & 3) - ml;
                        }
            }
            break;
        case T_Q3_K:        // low 2 bits in bit planes, the third bit in hmask (set = no -4 offset), sixteen 6-bit scales packed in 12 bytes
            for (uint64_t b = 0; b < k / 256; ++b) {
                const uint8_t* p = src + b * 110;
                const uint8_t* hm = p; const uint8_t* q = p + 32; const uint8_t* s12 = p + 96;
                uint16_t dh; memcpy(&dh, p + 108, 2);
                const float d = f16(dh);
                int scales[16];
                for (int j = 0; j < 16; ++j) {
                    const int lo = j < 8 ? (s12[j] & 0xF) : (s12[j - 8] >> 4);
                    const int hi = (s12[8 + j % 4] >> (2 * (j / 4))) & 3;
                    scales[j] = (lo | (hi << 4)) - 32;
                }
                float* y = out + b * 256;
                int is = 0, bit = 0;
                for (int half = 0; half < 2; ++half, q += 32)
                    for (int sh = 0; sh < 8; sh += 2, ++bit)
                        for (int part = 0; part < 2; ++part, ++is) {
                            const float dl = d * scales[is];
                            for (int l = 0; l < 16; ++l) {
                                const int idx = part * 16 + l;
                                *y++ = dl * ((int)((q[idx] >> sh) & 3) - (((hm[idx] >> bit) & 1) ? 0 : 4));
                            }
                        }
            }
            break;
        default: throw std::runtime_error("decode_row: unsupported type");
    }
}

inline float dot(const float* a, const float* b, uint64_t n) {
    __m256 acc0 = _mm256_setzero_ps(), acc1 = _mm256_setzero_ps();
    uint64_t i = 0;
    for (; i + 16 <= n; i += 16) {
        acc0 = _mm256_fmadd_ps(_mm256_loadu_ps(a + i), _mm256_loadu_ps(b + i), acc0);
        acc1 = _mm256_fmadd_ps(_mm256_loadu_ps(a + i + 8), _mm256_loadu_ps(b + i + 8), acc1);
    }
    acc0 = _mm256_add_ps(acc0, acc1);
    __m128 s = _mm_add_ps(_mm256_castps256_ps128(acc0), _mm256_extractf128_ps(acc0, 1));
    s = _mm_hadd_ps(s, s); s = _mm_hadd_ps(s, s);
    float r = _mm_cvtss_f32(s);
    for (; i < n; ++i) r += a[i] * b[i];
    return r;
}

}  // namespace dlp_gguf

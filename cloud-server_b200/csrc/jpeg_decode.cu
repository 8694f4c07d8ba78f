// Baseline JPEG decode for the upload path: what cv2.imread(dir) (apps/preprocess/preprocess.py:28, :40) and
// PIL.Image.open (apps/construction/util/construct_distribute.py:446, construct_inference.py:312) return, bit for bit.
// Both wheels embed libjpeg-turbo; this file follows its baseline path:
//   host    marker parsing + Huffman decoding (ITU T.81 F.2.2; serial by nature, one host thread per picture)
//           -> per 8x8 block a 64-bit map of the nonzero coefficients (zigzag positions) + their values: the compressed
//              domain crosses PCIe, 3-10x fewer bytes than decoded pixels for photographic content
//   device  jpeg_idct_kernel   dequantise + jidctint.c "islow" inverse DCT (CONST_BITS 13, PASS1_BITS 2, exact integers),
//                              8 threads per block: a column each in pass 1, a row each in pass 2
//           jpeg_color_kernel  jdsample.c "fancy" (triangle) chroma upsampling for 2x2 / 2x1 sampling evaluated per output
//                              pixel from the chroma planes + jdcolor.c fixed-point YCbCr -> RGB, written as BGR / RGB / gray
// Progressive and arithmetic-coded files return NC_EUNSUPPORTED.
#include "nc_internal.h"
#include <atomic>
#include <thread>

namespace {

const unsigned char kZigzag[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
                                   35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// q95 pictures carry many large coefficients whose (run, size) codes are 10-16 bits long: with a 9-bit table a third of the symbols
// fell through to the bit-by-bit search
constexpr int kLook = 12;

struct HuffTab {
    bool present = false;
    unsigned char look_nbits[1 << kLook], look_sym[1 << kLook];   // lookahead table: codes of up to kLook bits in one step
    int maxcode[18];                                // largest code of each length (-1 if none), [17] = sentinel
    int valoffset[17];
    unsigned char vals[256];
    // AC tables: one load per coefficient where the code AND its value bits fit the lookahead.  Entry: bits 0-4 bits consumed,
    // bit 5 'symbol only' (bits 8-15 = run << 4 | size, value bits follow in the stream), bit 6 'value included' (bits 8-11 run,
    // bits 16-31 the extended coefficient); 0 = code longer than the lookahead
    unsigned fast[1 << kLook];
};

struct Comp { int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0; };

struct Header {
    int width = 0, height = 0, ncomp = 0;
    Comp comp[3];
    int qt[4][64];            // natural order
    bool qt_present[4] = {false, false, false, false};
    HuffTab dc[4], ac[4];
    int restart_interval = 0;
    const unsigned char* ecs = nullptr;
    size_t ecs_len = 0;
    int hmax = 1, vmax = 1;
};

bool build_hufftab(const unsigned char* counts, const unsigned char* syms, int nsym, HuffTab* t) {
    int code = 0, k = 0;
    int huffcode[257];
    unsigned char huffsize[257];
    for (int l = 1; l <= 16; ++l)
        for (int i = 0; i < counts[l - 1]; ++i) { if (k >= 256) return false; huffsize[k++] = (unsigned char)l; }
    if (k != nsym) return false;
    huffsize[k] = 0;
    k = 0;
    int si = huffsize[0];
    while (huffsize[k]) {
        while (huffsize[k] == si) huffcode[k++] = code++;
        if (code > (1 << si)) return false;
        code <<= 1; ++si;
    }
    int p = 0;
    for (int l = 1; l <= 16; ++l) {
        if (counts[l - 1]) { t->valoffset[l] = p - huffcode[p]; p += counts[l - 1]; t->maxcode[l] = huffcode[p - 1]; }
        else { t->maxcode[l] = -1; t->valoffset[l] = 0; }
    }
    t->maxcode[17] = 0xFFFFF;
    memset(t->look_nbits, 0, sizeof(t->look_nbits));
    p = 0;
    for (int l = 1; l <= kLook; ++l)
        for (int i = 0; i < counts[l - 1]; ++i, ++p) {
            const int look = huffcode[p] << (kLook - l);
            for (int c = 0; c < (1 << (kLook - l)); ++c) { t->look_nbits[look + c] = (unsigned char)l; t->look_sym[look + c] = syms[p]; }
        }
    memset(t->fast, 0, sizeof(t->fast));
    p = 0;
    for (int l = 1; l <= kLook; ++l)
        for (int i = 0; i < counts[l - 1]; ++i, ++p) {
            const int rs = syms[p], r = rs >> 4, sz = rs & 15;
            const int rest = kLook - l;
            for (int c = 0; c < (1 << rest); ++c) {
                unsigned e = (unsigned)l | 0x20u | ((unsigned)rs << 8);
                if (sz && sz <= rest) {
                    const int bits = c >> (rest - sz);
                    const int v = bits < (1 << (sz - 1)) ? bits - (1 << sz) + 1 : bits;
                    e = (unsigned)(l + sz) | 0x40u | ((unsigned)r << 8) | ((unsigned)(v & 0xFFFF) << 16);
                }
                t->fast[(huffcode[p] << rest) + c] = e;
            }
        }
    memcpy(t->vals, syms, nsym);
    t->present = true;
    return true;
}

// returns NC_OK or an error code; msg receives a short reason
int parse_header(const unsigned char* d, size_t n, Header* h, const char** msg) {
    *msg = "";
    if (n < 4 || d[0] != 0xFF || d[1] != 0xD8) { *msg = "not a JPEG (no SOI marker)"; return NC_EINVAL; }
    size_t i = 2;
    bool have_sof = false;
    while (i + 3 < n) {
        if (d[i] != 0xFF) { *msg = "marker expected"; return NC_EINVAL; }
        while (i < n && d[i] == 0xFF) ++i;
        if (i >= n) break;
        const int m = d[i++];
        if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;
        if (m == 0xD9) break;
        if (i + 1 >= n) break;
        const size_t len = ((size_t)d[i] << 8) | d[i + 1];
        if (len < 2 || i + len > n) { *msg = "truncated segment"; return NC_EINVAL; }
        const unsigned char* s = d + i + 2;
        const size_t sl = len - 2;
        if (m == 0xDB) {
            size_t j = 0;
            while (j < sl) {
                const int pq = s[j] >> 4, tq = s[j] & 15;
                ++j;
                if (tq > 3 || j + (pq ? 128 : 64) > sl) { *msg = "bad DQT"; return NC_EINVAL; }
                for (int k = 0; k < 64; ++k) {
                    h->qt[tq][kZigzag[k]] = pq ? ((s[j + 2 * k] << 8) | s[j + 2 * k + 1]) : s[j + k];
                }
                j += pq ? 128 : 64;
                h->qt_present[tq] = true;
            }
        } else if (m == 0xC0 || m == 0xC1) {
            if (sl < 6 || s[0] != 8) { *msg = "only 8-bit samples"; return NC_EUNSUPPORTED; }
            h->height = (s[1] << 8) | s[2]; h->width = (s[3] << 8) | s[4]; h->ncomp = s[5];
            if ((h->ncomp != 1 && h->ncomp != 3) || sl < (size_t)6 + 3 * h->ncomp) { *msg = "1 or 3 components only"; return NC_EUNSUPPORTED; }
            for (int c = 0; c < h->ncomp; ++c) {
                h->comp[c].id = s[6 + 3 * c]; h->comp[c].h = s[7 + 3 * c] >> 4; h->comp[c].v = s[7 + 3 * c] & 15; h->comp[c].tq = s[8 + 3 * c];
                if (h->comp[c].h < 1 || h->comp[c].h > 2 || h->comp[c].v < 1 || h->comp[c].v > 2 || h->comp[c].tq > 3) { *msg = "unsupported sampling factors"; return NC_EUNSUPPORTED; }
                if (h->comp[c].h > h->hmax) h->hmax = h->comp[c].h;
                if (h->comp[c].v > h->vmax) h->vmax = h->comp[c].v;
            }
            have_sof = true;
        } else if (m >= 0xC2 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
            *msg = "progressive / lossless / arithmetic JPEG"; return NC_EUNSUPPORTED;
        } else if (m == 0xC4) {
            size_t j = 0;
            while (j + 17 <= sl) {
                const int tc = s[j] >> 4, th = s[j] & 15;
                int nsym = 0;
                for (int k = 0; k < 16; ++k) nsym += s[j + 1 + k];
                if (th > 3 || tc > 1 || nsym > 256 || j + 17 + nsym > sl) { *msg = "bad DHT"; return NC_EINVAL; }
                if (!build_hufftab(s + j + 1, s + j + 17, nsym, tc ? &h->ac[th] : &h->dc[th])) { *msg = "bad Huffman table"; return NC_EINVAL; }
                j += 17 + nsym;
            }
        } else if (m == 0xDD) {
            if (sl >= 2) h->restart_interval = (s[0] << 8) | s[1];
        } else if (m == 0xDA) {
            if (!have_sof) { *msg = "SOS before SOF"; return NC_EINVAL; }
            const int ns = s[0];
            if (ns != h->ncomp || sl < (size_t)1 + 2 * ns) { *msg = "non-interleaved scans are not supported"; return NC_EUNSUPPORTED; }
            for (int k = 0; k < ns; ++k) {
                bool found = false;
                for (int c = 0; c < h->ncomp; ++c)
                    if (h->comp[c].id == s[1 + 2 * k]) { h->comp[c].td = s[2 + 2 * k] >> 4; h->comp[c].ta = s[2 + 2 * k] & 15; found = c == k; }
                if (!found) { *msg = "scan component order"; return NC_EUNSUPPORTED; }
            }
            for (int c = 0; c < h->ncomp; ++c) {
                if (h->comp[c].td > 3 || h->comp[c].ta > 3 || !h->dc[h->comp[c].td].present || !h->ac[h->comp[c].ta].present || !h->qt_present[h->comp[c].tq]) {
                    *msg = "missing table"; return NC_EINVAL;
                }
                // the device code handles chroma at full or half resolution of a luma that has the largest factors
                if (c > 0 && !((h->comp[c].h == h->hmax || h->comp[c].h * 2 == h->hmax) && (h->comp[c].v == h->vmax || (h->comp[c].v * 2 == h->vmax && h->comp[c].h * 2 == h->hmax)))) {
                    *msg = "unsupported chroma sampling"; return NC_EUNSUPPORTED;
                }
            }
            if (h->ncomp == 3 && (h->comp[0].h != h->hmax || h->comp[0].v != h->vmax || h->comp[1].h != h->comp[2].h || h->comp[1].v != h->comp[2].v)) {
                *msg = "unsupported sampling layout"; return NC_EUNSUPPORTED;
            }
            if (h->ncomp == 1) { h->comp[0].h = h->comp[0].v = 1; h->hmax = h->vmax = 1; }     // a single-component scan is not interleaved
            h->ecs = d + i + len; h->ecs_len = n - (i + len);
            if (h->width < 1 || h->height < 1) { *msg = "empty picture"; return NC_EINVAL; }
            return NC_OK;
        }
        i += len;
    }
    *msg = "no SOS marker";
    return NC_EINVAL;
}

struct BitReader {
    const unsigned char* p; const unsigned char* end;
    unsigned long long acc = 0; int nbits = 0;       // the low `nbits` bits of acc are unread data, oldest bit highest
    bool hit_marker = false;
    inline void fill_bytes() {
        while (nbits <= 56) {
            unsigned b = 0;
            if (!hit_marker && p < end) {
                b = *p;
                if (b == 0xFF) {
                    const unsigned nx = p + 1 < end ? p[1] : 0xD9;
                    if (nx == 0) p += 2;
                    else { hit_marker = true; b = 0; }      // a marker: zeros from here on, position kept
                } else ++p;
            }
            acc = (acc << 8) | b;
            nbits += 8;
        }
    }
    // at least 57 bits afterwards.  Fast path: the next 8 bytes hold no 0xFF (no stuffing, no marker): take as many whole bytes
    // as fit with one load and one byte swap
    inline void fill() {
        if (!hit_marker && p + 8 <= end) {
            unsigned long long v;
            memcpy(&v, p, 8);
            const unsigned long long x = ~v;
            if (!((x - 0x0101010101010101ull) & ~x & 0x8080808080808080ull)) {
                const int nbytes = (64 - nbits) >> 3;
                v = __builtin_bswap64(v);
                acc = nbytes == 8 ? v : ((acc << (8 * nbytes)) | (v >> (64 - 8 * nbytes)));
                p += nbytes; nbits += 8 * nbytes;
                return;
            }
        }
        fill_bytes();
    }
    // the *_nc forms assume the caller made sure enough bits are buffered (need32() covers one Huffman code + its value bits)
    inline void need32() { if (nbits < 32) fill(); }
    inline unsigned peek_nc(int k) const { return (unsigned)((acc >> (nbits - k)) & ((1u << k) - 1)); }
    inline unsigned get_nc(int k) { nbits -= k; return (unsigned)((acc >> nbits) & ((1u << k) - 1)); }
    inline int decode_nc(const HuffTab& t) {
        const unsigned look = peek_nc(kLook);
        int nb = t.look_nbits[look];
        if (nb) { nbits -= nb; return t.look_sym[look]; }
        nb = kLook + 1;
        int code = (int)peek_nc(nb);
        while (code > t.maxcode[nb]) { ++nb; if (nb > 16) return -1; code = (int)peek_nc(nb); }
        nbits -= nb;
        const int idx = code + t.valoffset[nb];
        return (idx >= 0 && idx < 256) ? t.vals[idx] : -1;
    }
    // byte-align, then move behind the next RSTn marker
    inline void restart() {
        acc = 0; nbits = 0; hit_marker = false;
        while (p + 1 < end && !(p[0] == 0xFF && p[1] >= 0xD0 && p[1] <= 0xD7)) ++p;
        if (p + 1 < end) p += 2;
    }
};

inline int extend(unsigned v, int s) { return (int)v < (1 << (s - 1)) ? (int)v - (1 << s) + 1 : (int)v; }

struct Decoded {       // one picture in the compressed-domain layout the device expands
    int ncomp = 0, hs = 1, vs = 1;               // chroma subsampling: hmax / h_chroma, vmax / v_chroma
    int bw[3] = {0, 0, 0}, bh[3] = {0, 0, 0};    // blocks per row / column of each component plane (MCU padded)
    int qt[3][64];
    std::vector<unsigned long long> mask;        // per block, component planes back to back, blocks row-major
    std::vector<unsigned> off;                   // per block: index of its first value
    std::vector<short> vals;
    int rc = NC_OK; const char* msg = "";
};

void decode_one(const unsigned char* d, size_t n, int want_w, int want_h, Decoded* out) {
    Header h;
    out->rc = parse_header(d, n, &h, &out->msg);
    if (out->rc != NC_OK) return;
    if (h.width != want_w || h.height != want_h) { out->rc = NC_EINVAL; out->msg = "picture size differs from the batch's width x height"; return; }
    const int mcux = (h.width + 8 * h.hmax - 1) / (8 * h.hmax), mcuy = (h.height + 8 * h.vmax - 1) / (8 * h.vmax);
    out->ncomp = h.ncomp;
    out->hs = h.ncomp == 3 ? h.hmax / h.comp[1].h : 1;
    out->vs = h.ncomp == 3 ? h.vmax / h.comp[1].v : 1;
    size_t nblocks = 0, base[3] = {0, 0, 0};
    for (int c = 0; c < h.ncomp; ++c) {
        out->bw[c] = mcux * h.comp[c].h; out->bh[c] = mcuy * h.comp[c].v;
        base[c] = nblocks;
        nblocks += (size_t)out->bw[c] * out->bh[c];
        memcpy(out->qt[c], h.qt[h.comp[c].tq], sizeof(int) * 64);
    }
    out->mask.assign(nblocks, 0ull);
    out->off.assign(nblocks, 0u);
    // decode into a per-thread scratch of the worst-case size (no capacity check per value, no zero fill per picture), then keep
    // exactly what was decoded
    static thread_local std::vector<short> scratch;
    if (scratch.size() < nblocks * 64) scratch.resize(nblocks * 64);
    short* const v0 = scratch.data();
    short* vp = v0;
    BitReader br{h.ecs, h.ecs + h.ecs_len};
    int pred[3] = {0, 0, 0};
    long long cnt = 0;
#define NC_JPEG_FAIL(m_) do { out->rc = NC_EINVAL; out->msg = m_; return; } while (0)
    // values are appended in DECODE order (MCU interleaved); every block records where its own run starts
    for (int my = 0; my < mcuy; ++my) {
        for (int mx = 0; mx < mcux; ++mx) {
            if (h.restart_interval && cnt && cnt % h.restart_interval == 0) { br.restart(); pred[0] = pred[1] = pred[2] = 0; }
            ++cnt;
            for (int c = 0; c < h.ncomp; ++c) {
                const HuffTab& dct = h.dc[h.comp[c].td];
                const HuffTab& act = h.ac[h.comp[c].ta];
                for (int by = 0; by < h.comp[c].v; ++by) {
                    for (int bx = 0; bx < h.comp[c].h; ++bx) {
                        const size_t b = base[c] + (size_t)(my * h.comp[c].v + by) * out->bw[c] + (mx * h.comp[c].h + bx);
                        unsigned long long m = 0;
                        out->off[b] = (unsigned)(vp - v0);
                        br.need32();                          // one code (<= 16 bits) + its value bits (<= 15)
                        int s = br.decode_nc(dct);
                        if (s < 0 || s > 15) NC_JPEG_FAIL("corrupt Huffman data (DC)");
                        if (s) pred[c] += extend(br.get_nc(s), s);
                        if (pred[c]) { m |= 1ull; *vp++ = (short)pred[c]; }
                        for (int k = 1; k < 64;) {
                            br.need32();
                            const unsigned e = act.fast[br.peek_nc(kLook)];
                            if (e & 0x40u) {                       // code and value bits in one table entry
                                br.nbits -= (int)(e & 31u);
                                k += (int)((e >> 8) & 15u);
                                if (k > 63) NC_JPEG_FAIL("corrupt Huffman data (run past the block)");
                                m |= 1ull << k;
                                *vp++ = (short)((int)e >> 16);
                                ++k;
                                continue;
                            }
                            int rs;
                            if (e & 0x20u) { br.nbits -= (int)(e & 31u); rs = (int)((e >> 8) & 255u); }
                            else rs = br.decode_nc(act);
                            if (rs < 0) NC_JPEG_FAIL("corrupt Huffman data (AC)");
                            const int r = rs >> 4;
                            s = rs & 15;
                            if (s == 0) { if (r == 15) { k += 16; continue; } break; }
                            k += r;
                            if (k > 63) NC_JPEG_FAIL("corrupt Huffman data (run past the block)");
                            m |= 1ull << k;
                            *vp++ = (short)extend(br.get_nc(s), s);
                            ++k;
                        }
                        out->mask[b] = m;
                    }
                }
            }
        }
    }
#undef NC_JPEG_FAIL
    out->vals.assign(v0, vp);
}

// ------------------------------------------------------------------------------------------------------------ device
struct ImgDesc {            // one per picture, in device memory
    long long blk0;         // first block of this picture in the batch's mask / off arrays
    long long val0;         // first value of this picture in the batch's value array
    long long plane0;       // byte offset of its component planes in the plane scratch
    int ncomp, hs, vs;
    int bw[3], bh[3];
    int qt[3][64];
};

__constant__ unsigned char c_zigzag[64];

__device__ __forceinline__ int descale(int x, int n) { return (x + (1 << (n - 1))) >> n; }

// jidctint.c jpeg_idct_islow, one 1-D pass on d[0..7]; SH = left shift of the even part, DS = descale of the outputs
template <int DS>
__device__ __forceinline__ void idct8(const int* d, int* o) {
    int z2 = d[2], z3 = d[6];
    int z1 = (z2 + z3) * 4433;
    const int tmp2 = z1 + z3 * (-15137);
    const int tmp3 = z1 + z2 * 6270;
    z2 = d[0]; z3 = d[4];
    const int tmp0 = (z2 + z3) << 13, tmp1 = (z2 - z3) << 13;
    const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
    int t0 = d[7], t1 = d[5], t2 = d[3], t3 = d[1];
    z1 = t0 + t3; z2 = t1 + t2; z3 = t0 + t2;
    int z4 = t1 + t3;
    const int z5 = (z3 + z4) * 9633;
    t0 *= 2446; t1 *= 16819; t2 *= 25172; t3 *= 12299;
    z1 *= -7373; z2 *= -20995; z3 *= -16069; z4 *= -3196;
    z3 += z5; z4 += z5;
    t0 += z1 + z3; t1 += z2 + z4; t2 += z2 + z3; t3 += z1 + z4;
    o[0] = descale(tmp10 + t3, DS); o[7] = descale(tmp10 - t3, DS);
    o[1] = descale(tmp11 + t2, DS); o[6] = descale(tmp11 - t2, DS);
    o[2] = descale(tmp12 + t1, DS); o[5] = descale(tmp12 - t1, DS);
    o[3] = descale(tmp13 + t0, DS); o[4] = descale(tmp13 - t0, DS);
}

__device__ __forceinline__ unsigned range_limit(int v) {      // IDCT_range_limit(cinfo)[v & RANGE_MASK]
    const int idx = v & 1023;
    return idx < 128 ? idx + 128 : idx < 512 ? 255 : idx < 896 ? 0 : idx - 896;
}

// grid.x: groups of 32 blocks, grid.y: picture.  8 threads per 8x8 block.
__global__ void __launch_bounds__(256)
jpeg_idct_kernel(const ImgDesc* __restrict__ descs, const unsigned long long* __restrict__ masks, const unsigned* __restrict__ offs,
                 const short* __restrict__ vals, unsigned char* __restrict__ planes) {
    __shared__ int ws[32][8][9];
    const ImgDesc& D = descs[blockIdx.y];
    const int lb = threadIdx.x >> 3, j = threadIdx.x & 7;
    long long nb = 0;
    for (int c = 0; c < D.ncomp; ++c) nb += (long long)D.bw[c] * D.bh[c];
    const long long b = (long long)blockIdx.x * 32 + lb;
    const bool live = b < nb;
    int comp = 0; long long bc = b, pl = D.plane0;
    if (live) {
        while (bc >= (long long)D.bw[comp] * D.bh[comp]) { bc -= (long long)D.bw[comp] * D.bh[comp]; pl += (long long)D.bw[comp] * D.bh[comp] * 64; ++comp; }
        // expand: thread j owns zigzag positions 8j .. 8j+7
#pragma unroll
        for (int r = 0; r < 8; ++r) ws[lb][r][j] = 0;
    }
    __syncwarp();
    if (live) {
        const unsigned long long m = masks[D.blk0 + b];
        const unsigned byte = (unsigned)(m >> (8 * j)) & 0xFFu;
        const short* v = vals + D.val0 + offs[D.blk0 + b] + __popcll(m & ((1ull << (8 * j)) - 1ull));
        int t = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            if (byte & (1u << k)) {
                const int nat = c_zigzag[8 * j + k];
                ws[lb][nat >> 3][nat & 7] = (int)v[t++] * D.qt[comp][nat];
            }
        }
    }
    __syncwarp();
    int in[8], o[8];
    if (live) {
#pragma unroll
        for (int r = 0; r < 8; ++r) in[r] = ws[lb][r][j];         // pass 1: column j
        idct8<13 - 2>(in, o);
    }
    __syncwarp();
    if (live) {
#pragma unroll
        for (int r = 0; r < 8; ++r) ws[lb][r][j] = o[r];
    }
    __syncwarp();
    if (live) {
#pragma unroll
        for (int c = 0; c < 8; ++c) in[c] = ws[lb][j][c];         // pass 2: row j
        idct8<13 + 2 + 3>(in, o);
        const int bx = (int)(bc % D.bw[comp]), by = (int)(bc / D.bw[comp]);
        const unsigned lo = range_limit(o[0]) | (range_limit(o[1]) << 8) | (range_limit(o[2]) << 16) | (range_limit(o[3]) << 24);
        const unsigned hi = range_limit(o[4]) | (range_limit(o[5]) << 8) | (range_limit(o[6]) << 16) | (range_limit(o[7]) << 24);
        unsigned char* row = planes + pl + ((long long)(by * 8 + j) * D.bw[comp] + bx) * 8;
        *reinterpret_cast<uint2*>(row) = make_uint2(lo, hi);
    }
}

__device__ __forceinline__ int chroma_at(const unsigned char* __restrict__ p, int pitch, int dw, int dh, int hs, int vs, int x, int y) {
    if (hs == 1) return p[(long long)y * pitch + x];
    const int cx = x >> 1;
    if (vs == 1) {                                          // h2v1_fancy_upsample
        const int t = p[(long long)y * pitch + cx];
        if (x & 1) return cx == dw - 1 ? t : (3 * t + p[(long long)y * pitch + cx + 1] + 2) >> 2;
        return cx == 0 ? t : (3 * t + p[(long long)y * pitch + cx - 1] + 1) >> 2;
    }
    const int cy = y >> 1;                                  // h2v2_fancy_upsample
    const int oy = (y & 1) ? min(cy + 1, dh - 1) : max(cy - 1, 0);
    const unsigned char* r0 = p + (long long)cy * pitch;
    const unsigned char* r1 = p + (long long)oy * pitch;
    const int cs = 3 * r0[cx] + r1[cx];
    if (x & 1) return cx == dw - 1 ? (cs * 4 + 7) >> 4 : (cs * 3 + 3 * r0[cx + 1] + r1[cx + 1] + 7) >> 4;
    return cx == 0 ? (cs * 4 + 8) >> 4 : (cs * 3 + 3 * r0[cx - 1] + r1[cx - 1] + 8) >> 4;
}

// grid.y: picture; one thread per output pixel
__global__ void __launch_bounds__(256)
jpeg_color_kernel(const ImgDesc* __restrict__ descs, const unsigned char* __restrict__ planes, unsigned char* __restrict__ out,
                  long long img_stride, int width, int height, int out_channels, int bgr) {
    const ImgDesc& D = descs[blockIdx.y];
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)width * height) return;
    const int x = (int)(idx % width), y = (int)(idx / width);
    const unsigned char* py = planes + D.plane0;
    const int yv = py[(long long)y * D.bw[0] * 8 + x];
    unsigned char* o = out + blockIdx.y * img_stride + idx * out_channels;
    if (D.ncomp == 1) {
        o[0] = (unsigned char)yv;
        if (out_channels == 3) { o[1] = (unsigned char)yv; o[2] = (unsigned char)yv; }
        return;
    }
    const unsigned char* pcb = py + (long long)D.bw[0] * D.bh[0] * 64;
    const unsigned char* pcr = pcb + (long long)D.bw[1] * D.bh[1] * 64;
    const int dw = (width + D.hs - 1) / D.hs, dh = (height + D.vs - 1) / D.vs;
    const int cb = chroma_at(pcb, D.bw[1] * 8, dw, dh, D.hs, D.vs, x, y) - 128;
    const int cr = chroma_at(pcr, D.bw[2] * 8, dw, dh, D.hs, D.vs, x, y) - 128;
    const int r = min(max(yv + ((91881 * cr + 32768) >> 16), 0), 255);
    const int g = min(max(yv + ((-22554 * cb + 32768 - 46802 * cr) >> 16), 0), 255);
    const int b = min(max(yv + ((116130 * cb + 32768) >> 16), 0), 255);
    if (out_channels == 3) { o[0] = (unsigned char)(bgr ? b : r); o[1] = (unsigned char)g; o[2] = (unsigned char)(bgr ? r : b); }
    else o[0] = (unsigned char)yv;        // (callers wanting PIL's / cv2's gray conversion take 3 channels and convert)
}

bool g_zigzag_uploaded[64] = {};

}  // namespace

extern "C" int nc_jpeg_info(const uint8_t* host_jpeg, size_t size, int* width, int* height, int* components, int* hsamp, int* vsamp) {
    if (!host_jpeg) return nc_set_error(nullptr, NC_EINVAL, "host_jpeg is NULL");
    Header h;
    const char* msg;
    const int rc = parse_header(host_jpeg, size, &h, &msg);
    if (rc != NC_OK) return nc_set_error(nullptr, rc, "JPEG header: %s", msg);
    if (width) *width = h.width;
    if (height) *height = h.height;
    if (components) *components = h.ncomp;
    if (hsamp) *hsamp = h.ncomp == 3 ? h.hmax / h.comp[1].h : 1;
    if (vsamp) *vsamp = h.ncomp == 3 ? h.vmax / h.comp[1].v : 1;
    return NC_OK;
}

// Host only: the Huffman stage alone, expanded to dense blocks -- what the device kernels are fed, for checking that stage
// without a GPU.  coefs[b][64] (natural order, quantised), component planes back to back, blocks row-major, MCU padded.
extern "C" int nc_jpeg_entropy_decode(const uint8_t* host_jpeg, size_t size, int16_t* coefs, int64_t capacity_blocks, int64_t* nblocks) {
    if (!host_jpeg || !nblocks) return nc_set_error(nullptr, NC_EINVAL, "NULL argument");
    Header h;
    const char* msg;
    int rc = parse_header(host_jpeg, size, &h, &msg);
    if (rc != NC_OK) return nc_set_error(nullptr, rc, "JPEG header: %s", msg);
    Decoded d;
    decode_one(host_jpeg, size, h.width, h.height, &d);
    if (d.rc != NC_OK) return nc_set_error(nullptr, d.rc, "JPEG: %s", d.msg);
    *nblocks = (int64_t)d.mask.size();
    if (!coefs) return NC_OK;
    if (capacity_blocks < *nblocks) return nc_set_error(nullptr, NC_EINVAL, "capacity %lld < %lld blocks", (long long)capacity_blocks, (long long)*nblocks);
    memset(coefs, 0, (size_t)*nblocks * 64 * sizeof(int16_t));
    for (size_t b = 0; b < d.mask.size(); ++b) {
        const short* v = d.vals.data() + d.off[b];
        for (int k = 0; k < 64; ++k)
            if (d.mask[b] & (1ull << k)) coefs[b * 64 + kZigzag[k]] = *v++;
    }
    return NC_OK;
}

extern "C" int nc_jpeg_decode(nc_ctx* ctx, const uint8_t* const* host_jpegs, const size_t* sizes, int count, int width, int height,
                              int out_channels, int channel_order, uint8_t* out, int64_t img_stride, int host_threads,
                              int64_t* h2d_bytes, void* stream) {
    if (!ctx) return nc_set_error(nullptr, NC_EINVAL, "ctx is NULL");
    if (count < 0 || (count > 0 && (!host_jpegs || !sizes || !out))) return nc_set_error(ctx, NC_EINVAL, "bad arguments");
    if (out_channels != 1 && out_channels != 3) return nc_set_error(ctx, NC_EINVAL, "out_channels must be 1 or 3");
    if (width < 1 || height < 1 || img_stride < (int64_t)width * height * out_channels) return nc_set_error(ctx, NC_EINVAL, "bad output shape / stride");
    if (h2d_bytes) *h2d_bytes = 0;
    if (count == 0) return NC_OK;
    NC_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    if (ctx->device >= 0 && ctx->device < 64 && !g_zigzag_uploaded[ctx->device]) {
        NC_CUDA(ctx, cudaMemcpyToSymbol(c_zigzag, kZigzag, 64));
        g_zigzag_uploaded[ctx->device] = true;
    }
    // ---- host: Huffman decoding, one picture per task
    std::vector<Decoded> dec((size_t)count);
    int nthr = host_threads > 0 ? host_threads : (int)std::thread::hardware_concurrency();
    if (nthr < 1) nthr = 1;
    if (nthr > count) nthr = count;
    std::atomic<int> next{0};
    auto work = [&]() {
        for (;;) {
            const int i = next.fetch_add(1);
            if (i >= count) return;
            if (!host_jpegs[i]) { dec[i].rc = NC_EINVAL; dec[i].msg = "NULL picture"; continue; }
            decode_one(host_jpegs[i], sizes[i], width, height, &dec[i]);
        }
    };
    if (nthr == 1) work();
    else {
        std::vector<std::thread> pool;
        for (int t = 0; t < nthr; ++t) pool.emplace_back(work);
        for (auto& t : pool) t.join();
    }
    for (int i = 0; i < count; ++i)
        if (dec[i].rc != NC_OK) return nc_set_error(ctx, dec[i].rc, "picture %d: %s", i, dec[i].msg);
    // ---- pack into one pinned staging buffer: [descs][masks][offs][vals]
    std::vector<ImgDesc> descs((size_t)count);
    long long nblk = 0, nval = 0, nplane = 0, max_blocks = 0;
    for (int i = 0; i < count; ++i) {
        ImgDesc& D = descs[i];
        D.blk0 = nblk; D.val0 = nval; D.plane0 = nplane;
        D.ncomp = dec[i].ncomp; D.hs = dec[i].hs; D.vs = dec[i].vs;
        long long b = 0;
        for (int c = 0; c < 3; ++c) {
            D.bw[c] = dec[i].bw[c]; D.bh[c] = dec[i].bh[c];
            b += (long long)D.bw[c] * D.bh[c];
            memcpy(D.qt[c], dec[i].qt[c], sizeof(int) * 64);
        }
        nblk += b; nval += (long long)dec[i].vals.size(); nplane += b * 64;
        if (b > max_blocks) max_blocks = b;
    }
    const size_t sz_desc = (size_t)count * sizeof(ImgDesc), sz_mask = (size_t)nblk * 8, sz_off = (size_t)nblk * 4;
    const size_t sz_val = ((size_t)nval * 2 + 15) & ~(size_t)15;
    const size_t o_mask = (sz_desc + 15) & ~(size_t)15, o_off = o_mask + sz_mask, o_val = (o_off + sz_off + 15) & ~(size_t)15;
    const size_t total = o_val + sz_val;
    int rc = nc_pinned_reserve(ctx, total);
    if (rc != NC_OK) return rc;
    // the staging buffer may still feed the previous call's copy on this stream
    NC_CUDA(ctx, cudaStreamSynchronize(st));
    unsigned char* hp = static_cast<unsigned char*>(ctx->pinned);
    memcpy(hp, descs.data(), sz_desc);
    {
        // packing is ~0.6 MB per picture: spread it over the same host threads
        std::atomic<int> nextp{0};
        auto pack = [&]() {
            for (;;) {
                const int i = nextp.fetch_add(1);
                if (i >= count) return;
                memcpy(hp + o_mask + (size_t)descs[i].blk0 * 8, dec[i].mask.data(), dec[i].mask.size() * 8);
                memcpy(hp + o_off + (size_t)descs[i].blk0 * 4, dec[i].off.data(), dec[i].off.size() * 4);
                memcpy(hp + o_val + (size_t)descs[i].val0 * 2, dec[i].vals.data(), dec[i].vals.size() * 2);
            }
        };
        if (nthr == 1) pack();
        else {
            std::vector<std::thread> pool;
            for (int t = 0; t < nthr; ++t) pool.emplace_back(pack);
            for (auto& t : pool) t.join();
        }
    }
    const size_t plane_off = (total + 255) & ~(size_t)255;
    rc = nc_ws_reserve(ctx, plane_off + (size_t)nplane);
    if (rc != NC_OK) return rc;
    unsigned char* dp = static_cast<unsigned char*>(ctx->ws);
    NC_CUDA(ctx, cudaMemcpyAsync(dp, hp, total, cudaMemcpyHostToDevice, st));
    if (h2d_bytes) *h2d_bytes = (int64_t)total;
    const ImgDesc* d_desc = reinterpret_cast<const ImgDesc*>(dp);
    dim3 g1((unsigned)((max_blocks + 31) / 32), (unsigned)count);
    jpeg_idct_kernel<<<g1, 256, 0, st>>>(d_desc, reinterpret_cast<const unsigned long long*>(dp + o_mask), reinterpret_cast<const unsigned*>(dp + o_off),
                                         reinterpret_cast<const short*>(dp + o_val), dp + plane_off);
    NC_LAUNCH_CHECK(ctx);
    dim3 g2((unsigned)(((long long)width * height + 255) / 256), (unsigned)count);
    jpeg_color_kernel<<<g2, 256, 0, st>>>(d_desc, dp + plane_off, out, img_stride, width, height, out_channels, channel_order == 1 ? 1 : 0);
    NC_LAUNCH_CHECK(ctx);
    return NC_OK;
}

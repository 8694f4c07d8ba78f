// Host-only helpers behind the C ABI (no GPU involved).
//
// nc_crc32c: CRC-32C (Castagnoli, reflected polynomial 0x82F63B78) as TensorFlow's checkpoint "V2" bundle stores it per
// tensor and per table block (masked) -- needed to read and write the files tf.train.Saver produces for the reference
// (apps/construction/util/construct_distribute.py:375, :385-392; construct_inference.py:336-339).  Slicing-by-8.
#include "nc_internal.h"

namespace {
struct Crc32cTables {
    uint32_t t[8][256];
    Crc32cTables() {
        for (uint32_t i = 0; i < 256; ++i) {
            uint32_t c = i;
            for (int k = 0; k < 8; ++k) c = (c & 1u) ? (c >> 1) ^ 0x82F63B78u : c >> 1;
            t[0][i] = c;
        }
        for (uint32_t i = 0; i < 256; ++i)
            for (int s = 1; s < 8; ++s) t[s][i] = (t[s - 1][i] >> 8) ^ t[0][t[s - 1][i] & 0xFFu];
    }
};
const Crc32cTables g_crc;
}  // namespace

extern "C" uint32_t nc_crc32c(const void* data, size_t n, uint32_t crc) {
    const unsigned char* p = static_cast<const unsigned char*>(data);
    uint32_t c = ~crc;
    while (n && (reinterpret_cast<uintptr_t>(p) & 7u)) { c = (c >> 8) ^ g_crc.t[0][(c ^ *p++) & 0xFFu]; --n; }
    while (n >= 8) {
        uint64_t w;
        memcpy(&w, p, 8);
        w ^= c;
        c = g_crc.t[7][w & 0xFF] ^ g_crc.t[6][(w >> 8) & 0xFF] ^ g_crc.t[5][(w >> 16) & 0xFF] ^ g_crc.t[4][(w >> 24) & 0xFF] ^
            g_crc.t[3][(w >> 32) & 0xFF] ^ g_crc.t[2][(w >> 40) & 0xFF] ^ g_crc.t[1][(w >> 48) & 0xFF] ^ g_crc.t[0][(w >> 56) & 0xFF];
        p += 8; n -= 8;
    }
    while (n--) c = (c >> 8) ^ g_crc.t[0][(c ^ *p++) & 0xFFu];
    return ~c;
}

// Host side of the COCO RLE wire format: the compressed string <-> counts codec of pycocotools
// (cocoapi/common/maskApi.c: rleToString / rleFrString — "similar to LEB128 but using 6 bits/char and ascii chars
// 48-111", counts from the fourth on stored as the difference to the count two places earlier).  The reference keeps
// every dataset mask in this form (`rle['counts'].decode('ascii')`, tools/ours/panoptic2eee.py:61-64;
// maskrefiner/data/datasets/register_uoais_sim_panoptic.py:121-144) and hands it to mask_util.decode
// (maskrefiner/data/dataset_mappers/perturbed_panoptic_dataset_mapper.py:195-214).  A string is a few hundred bytes per
// mask: this stays on the host, the run lengths go to quber_rle_decode on the device.
#include <stddef.h>
#include <stdint.h>

#include "../../include/quber_b200.h"

extern "C" long long quber_rle_from_string(const char *s, size_t len, uint32_t *counts, size_t cap) {
    if (!s || (!counts && cap)) return QUBER_E_INVALID;
    size_t m = 0, p = 0;
    while (p < len && s[p]) {
        long long x = 0;
        int k = 0, more = 1;
        while (more) {
            if (p >= len || !s[p]) return QUBER_E_INVALID;             // truncated inside a count
            const int c = (int)s[p] - 48;
            if (c < 0 || c > 63 || k > 12) return QUBER_E_INVALID;     // outside the alphabet / longer than any 64-bit value
            x |= (long long)(c & 0x1f) << (5 * k);
            more = c & 0x20;
            ++p; ++k;
            if (!more && (c & 0x10)) x |= (long long)(~0ULL << (5 * k)); // sign extension
        }
        if (m > 2 && m - 2 < cap) x += (long long)counts[m - 2];
        if (m < cap) counts[m] = (uint32_t)x;
        ++m;
    }
    return (long long)m;                                               // > cap: the caller's buffer was too small
}

extern "C" long long quber_rle_to_string(const uint32_t *counts, size_t m, char *s, size_t cap) {
    if ((!counts && m) || !s || cap == 0) return QUBER_E_INVALID;
    size_t p = 0;
    for (size_t i = 0; i < m; ++i) {
        long long x = (long long)counts[i];
        if (i > 2) x -= (long long)counts[i - 2];
        int more = 1;
        while (more) {
            int c = (int)(x & 0x1f);
            x >>= 5;
            more = (c & 0x10) ? x != -1 : x != 0;
            if (more) c |= 0x20;
            if (p + 1 < cap) s[p] = (char)(c + 48);
            ++p;
        }
    }
    s[p < cap ? p : cap - 1] = 0;
    return (long long)p;                                               // >= cap: the string did not fit (needs p + 1 bytes)
}

/* A plain C11 caller of include/vsp_b200.h (built and run by tests/test_abi.py, no GPU needed): the header is C, the
 * library links from C, the host-only entry points work, and a context without a CUDA device is an error code — the
 * binding shape INTEGRATION.md gives for cgo / JNI / ctypes hosts. */
#include <stdio.h>
#include <string.h>

#include "vsp_b200.h"

static int fail(const char *what) {
    printf("FAIL %s\n", what);
    return 1;
}

int main(void) {
    /* serialize_bitarray: bit i in octet i / 8, most significant bit first */
    uint8_t bits[12] = {1, 0, 0, 0, 0, 0, 0, 1, 0, 1, 0, 0}, octets[2], back[12];
    if (vsp_codec_bitarray_write(bits, 12, octets) != VSP_OK || octets[0] != 0x81 || octets[1] != 0x40) return fail("bitarray_write");
    if (vsp_codec_bitarray_read(octets, 2, 12, back) != VSP_OK || memcmp(bits, back, 12)) return fail("bitarray_read");
    if (vsp_codec_bitarray_read(octets, 1, 12, back) != VSP_E_INVALID) return fail("bitarray_read length check");
    /* the point at infinity travels as 0xc0 00.. and back */
    uint64_t inf[12] = {0}, pt[12];
    uint8_t enc[48];
    if (vsp_codec_g1_compress(inf, enc) != VSP_OK || enc[0] != 0xc0) return fail("g1_compress(infinity)");
    memset(pt, 0xff, sizeof(pt));
    if (vsp_codec_g1_decompress(enc, pt) != VSP_OK || memcmp(pt, inf, sizeof(pt))) return fail("g1_decompress(infinity)");
    enc[0] = 0x40; /* C bit clear */
    if (vsp_codec_g1_decompress(enc, pt) != VSP_E_INVALID) return fail("g1_decompress rejects an uncompressed flag");
    /* make_evaluation_domain's choice for the ballot circuit's size */
    int kind = -1;
    uint64_t m = 0;
    if (vsp_domain_plan(31926 + 30 + 1, &kind, &m) != VSP_OK || m != 32768) return fail("domain_plan");
    /* no device, no context: an error code, never a CPU fallback */
    vsp_ctx *ctx = NULL;
    const int rc = vsp_ctx_create(0, NULL, &ctx);
    if (rc == VSP_OK) {
        vsp_ctx_destroy(ctx);
        printf("OK (a CUDA device is present) %s\n", vsp_version());
        return 0;
    }
    if (ctx != NULL) return fail("vsp_ctx_create left a handle behind on failure");
    printf("OK (no CUDA device: vsp_ctx_create = %d) %s\n", rc, vsp_version());
    return 0;
}

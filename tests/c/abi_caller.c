/* A plain C11 caller of the drop-in boundary (include/bitnet_b200.h): what a cgo / Rust-FFI binding does.
 * Exit code 0 = every check passed. With no GPU it checks the loud failure; with one, a real 2x2 matmul_i2s
 * (the identity case of crates/bitnet-kernels/src/cpu/fallback.rs:305-318). */
#include "bitnet_b200.h"

#include <stdio.h>
#include <string.h>

int main(void) {
    const int8_t a[4] = {1, 2, 3, 4};
    const uint8_t b[4] = {1, 0, 0, 1};
    float c[4] = {0, 0, 0, 0};
    /* argument validation happens before any device work (ffi/cpp_bridge.cpp:62-75) */
    if (bitnet_cpp_matmul_i2s(NULL, b, c, 2, 2, 2) != -1) return 10;
    if (!strstr(bitnet_cpp_get_last_error(), "Null pointer passed to matmul_i2s")) return 11;
    if (bitnet_cpp_matmul_i2s(a, b, c, 0, 2, 2) != -1) return 12;
    if (!strstr(bitnet_cpp_get_last_error(), "Invalid matrix dimensions")) return 13;
    if (!strstr(bitnet_b200_version(), "sm_100a")) return 14;
    if (!bitnet_cpp_is_available()) {
        /* no sm_100 device: init must FAIL (no CPU fallback) with a message */
        if (bitnet_cpp_init() == 0) return 20;
        if (strlen(bitnet_cpp_get_last_error()) == 0) return 21;
        puts("no device: loud failure ok");
        return 0;
    }
    if (bitnet_cpp_init() != 0) return 30;
    if (bitnet_cpp_matmul_i2s(a, b, c, 2, 2, 2) != 0) {
        fprintf(stderr, "%s\n", bitnet_cpp_get_last_error());
        return 31;
    }
    if (c[0] != 1.f || c[1] != 2.f || c[2] != 3.f || c[3] != 4.f) return 32;
    bitnet_cpp_cleanup();
    puts("device: matmul_i2s ok");
    return 0;
}

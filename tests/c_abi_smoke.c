/* Compiled and linked as plain C11 by tests/test_capi_symbols.py: proves that include/activesplat_b200.h is a valid
 * C header (no C++ leaks) and that every declared entry point links against libactivesplat_b200.so.
 * The functions called here are host-only helpers; nothing is launched. */
#include <stdio.h>
#include <string.h>

#include "activesplat_b200.h"

static char* never_called(void* ctx, size_t n) { (void)ctx; (void)n; return NULL; }

int main(void) {
    AsForwardArgs f;
    AsBackwardArgs b;
    AsFrameArgs fr;
    AsFrameBackwardArgs fb;
    AsScratchLayout lay;
    memset(&f, 0, sizeof f); memset(&b, 0, sizeof b); memset(&fr, 0, sizeof fr); memset(&fb, 0, sizeof fb);
    /* take the address of every compute entry point so the linker must resolve it */
    void* fns[] = {(void*)as_forward, (void*)as_backward, (void*)as_mark_visible, (void*)as_frame_forward,
                   (void*)as_frame_backward, (void*)as_tracking_loss, (void*)as_mapping_loss, (void*)as_adam_step,
                   (void*)as_sort_pairs, (void*)as_inclusive_scan_u32, (void*)as_inclusive_scan_u32_static,
                   (void*)as_profile_enable, (void*)as_profile_read, (void*)as_launch_count, (void*)never_called};
    if (as_scratch_layout(1000, 640, 480, 2500, 3, &lay) != AS_OK) return 1;
    if (as_forward(NULL, never_called, never_called, never_called, NULL, NULL) != AS_ERR_INVALID_ARG) return 2;
    if (as_higher_msb(1200) != 11) return 3;
    if (as_geometry_bytes(1000) == 0 || as_image_bytes(640, 480) == 0 || as_binning_bytes(2500, 6) == 0) return 4;
    if (as_static_scan_max_n() != 4194304) return 5;
    printf("%s: %zu entry points, geometry(1000)=%zu bytes, %s\n", as_version(), sizeof fns / sizeof fns[0],
           as_geometry_bytes(1000), as_error_string(AS_ERR_NO_COLORS));
    return 0;
}

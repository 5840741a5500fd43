/* abi_check.c — plain C11 translation unit over the three public headers: proves they are valid C (the language a
 * cgo / JNI / Rust-bindgen consumer parses them as), that libforma_rt.so links from compiled code without torch or
 * Python, and exercises the device-free part of the ABI on a FORMA_RT_HOST_ONLY handle.  Prints what it read back;
 * tests/test_cxx_host.py compares that with the Python mirror.  Every render entry point must fail with
 * FORMA_RT_ECUDA on such a handle (no CPU fallback). */
#include <stdio.h>
#include <string.h>

#include "forma_frag.h"
#include "forma_fx.h"
#include "forma_rt.h"

int main(void) {
    forma_rt_options opts;
    memset(&opts, 0, sizeof opts);
    opts.struct_size = sizeof opts;
    opts.flags = FORMA_RT_HOST_ONLY;
    forma_rt* m = NULL;
    if (forma_rt_create(&opts, 64, 48, &m) != FORMA_RT_OK) { fprintf(stderr, "create: %s\n", forma_rt_last_error()); return 1; }
    printf("abi %u\n", forma_rt_abi_version());
    for (int which = 0; which < 3; ++which) {
        size_t n = 0;
        if (forma_rt_load_builtin(m, which) || forma_rt_get_objects(m, NULL, 0, &n)) return 1;
        printf("builtin %d: %zu objects\n", which, n);
    }
    const float delta[3] = {0.0f, 0.0f, 0.016f};
    forma_rt_camera cam;
    if (forma_rt_camera_orbit(m, delta) || forma_rt_camera_get(m, &cam)) return 1;
    printf("camera %.9g %.9g %.9g radius %.9g rotation %.9g\n", cam.position[0], cam.position[1], cam.position[2], cam.radius, cam.rotation);
    unsigned char px[64 * 48 * 3];
    const int rc = forma_rt_update(m, 0, 0.016f, px);
    printf("update on a host-only handle: %d (%s)\n", rc, rc == FORMA_RT_ECUDA ? "FORMA_RT_ECUDA" : "unexpected");
    printf("effect 6 = %s\n", forma_fx_effect_name(FORMA_FX_FLIP_AXIS));
    float l[3];
    if (forma_frag_light_direction(3.0f, l)) return 1;
    printf("light %.9g %.9g %.9g\n", l[0], l[1], l[2]);
    forma_rt_destroy(m);
    return rc == FORMA_RT_ECUDA ? 0 : 1;
}

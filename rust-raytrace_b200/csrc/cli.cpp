// raytrace_b200 -- the reference's binary (src/main.rs:13-60) over the C ABI of include/raytrace_b200.h.
//
// Same behaviour for a user of `cargo run`: reads "test_scene.txt" (src/main.rs:16); an open / read / syntax error is ONE
// line "error: <message>" on stdout and a normal return, exit status 0 (src/main.rs:18,23,28: `println!("error: {}", e);
// return`), the message being the io::Error's or the SyntaxError's Display ("row:col: message"); warnings are printed as
// the parser meets them; the frame is rendered at the scene's Options and written to "out.bmp" (src/main.rs:34).  Everything between parsing and the file is the B200 path: this file only calls
// rt_scene_parse_file / rt_render / rt_bmp_write.  Positional arguments and flags are additions; with none given
// the two file names are upstream's.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/raytrace_b200.h"

static void usage() {
    std::fprintf(stderr,
                 "usage: raytrace_b200 [scene.txt [out.bmp]] [--width W] [--height H] [--antialias N] [--depth D]\n"
                 "                     [--seed S] [--accel auto|linear|bvh] [--fp strict|fast] [--device K] [--stats]\n"
                 "defaults: test_scene.txt -> out.bmp at the scene's own options (src/main.rs:16,34)\n");
}

int main(int argc, char** argv) {
    std::string scene_path = "test_scene.txt", out_path = "out.bmp";
    rt_render_opts o;
    rt_render_opts_default(&o);
    bool stats = false;
    int positional = 0;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto need = [&](const char* what) -> const char* {
            if (i + 1 >= argc) { std::fprintf(stderr, "%s needs a value\n", what); usage(); std::exit(2); }
            return argv[++i];
        };
        if (a == "--help" || a == "-h") { usage(); return 0; }
        else if (a == "--width") o.width = (uint32_t)std::strtoul(need("--width"), nullptr, 10);
        else if (a == "--height") o.height = (uint32_t)std::strtoul(need("--height"), nullptr, 10);
        else if (a == "--antialias") o.antialias = (uint32_t)std::strtoul(need("--antialias"), nullptr, 10);
        else if (a == "--depth") o.max_depth = (uint32_t)std::strtoul(need("--depth"), nullptr, 10);
        else if (a == "--seed") o.seed = std::strtoull(need("--seed"), nullptr, 10);
        else if (a == "--device") o.device = (int32_t)std::strtol(need("--device"), nullptr, 10);
        else if (a == "--stats") stats = true;
        else if (a == "--accel") {
            const std::string v = need("--accel");
            o.accel = v == "linear" ? RT_ACCEL_LINEAR : v == "bvh" ? RT_ACCEL_BVH : RT_ACCEL_AUTO;
        } else if (a == "--fp") {
            const std::string v = need("--fp");
            o.fp_mode = v == "fast" ? RT_FP_FAST : RT_FP_STRICT;
        } else if (!a.empty() && a[0] == '-') { std::fprintf(stderr, "unknown flag %s\n", a.c_str()); usage(); return 2; }
        else if (positional == 0) { scene_path = a; positional++; }
        else if (positional == 1) { out_path = a; positional++; }
        else { usage(); return 2; }
    }

    rt_scene* scene = nullptr;
    char err[512] = {0};
    int rc = rt_scene_parse_file(scene_path.c_str(), &scene, err, sizeof err);
    if (rc != RT_OK) {
        // src/main.rs:18,23,28: println!("error: {}", e); return  -- one line on stdout, exit status 0
        std::printf("error: %s\n", err[0] ? err : rt_last_error());
        return 0;
    }
    const char* warnings = rt_scene_warnings(scene);
    if (warnings && warnings[0]) std::printf("%s", warnings);  // upstream println!s them while parsing (src/serialize.rs:452-464)

    uint32_t w = 0, h = 0, aa = 0;
    rt_scene_get_options(scene, &w, &h, &aa);
    if (o.width) w = o.width;
    if (o.height) h = o.height;
    std::vector<uint8_t> rows((size_t)rt_bmp_bytewidth(w) * h);
    rt_stats st;
    std::memset(&st, 0, sizeof st);
    rc = rt_render(scene, &o, rows.data(), nullptr, nullptr, &st);
    if (rc != RT_OK) {
        std::fprintf(stderr, "render failed (%d): %s\n", rc, rt_last_error());
        rt_scene_free(scene);
        return 1;
    }
    rc = rt_bmp_write(out_path.c_str(), w, h, rows.data());
    if (rc != RT_OK) {
        std::fprintf(stderr, "cannot write %s: %s\n", out_path.c_str(), rt_last_error());
        rt_scene_free(scene);
        return 1;
    }
    if (stats)
        std::fprintf(stderr, "%ux%u: %llu rays (%llu primary, %llu secondary, %llu shadow) in %.3f ms on the device = %.1f Mrays/s, %llu kernel launches\n", w, h,
                     (unsigned long long)st.rays_total, (unsigned long long)st.rays_primary, (unsigned long long)st.rays_secondary,
                     (unsigned long long)st.rays_shadow, st.render_ms, st.render_ms > 0 ? st.rays_total / st.render_ms / 1e3 : 0.0,
                     (unsigned long long)st.kernel_launches);
    rt_scene_free(scene);
    return 0;
}

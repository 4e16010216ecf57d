// cli.cpp — `tiny_renderer`: the reference's command line (/root/reference/src/main.rs:12-26: `-p <asset
// folder>`, `-s <pipeline>`, defaults assets/diablo + default, 800x800) driving the CUDA backend through the C
// ABI, with a headless render-to-TGA mode in place of the show-image window (/root/reference/src/app.rs:148-153,
// 216-218 needs a graphics stack: unavailable on this box).  The camera / light parametrisation is the app
// loop's: look_from = (sin a, 0, cos a), light = (sin b, 0, cos b), up = +y (app.rs:200-207).
#include "../../include/tinyr.h"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

static int die(const char* what) {
    fprintf(stderr, "tiny_renderer: %s: %s\n", what, tr_last_error());
    return 1;
}

int main(int argc, char** argv) {
    std::string asset_path = "assets/diablo", pipeline = "default", out = "out.tga", dump_z, dump_shadow, replay, dump_visibility;
    uint32_t n_views = 0, n_lights = 0, n_gpus = 1, ring = 0;
    uint64_t seed = 0x5EED0A0ull;
    float fixed_dt = -1.0f;
    uint32_t width = 800, height = 800, frames = 1;
    float camera_angle = 0.0f, light_angle = 0.0f, camera_step = 0.0f;
    int device = 0;
    bool print_fps = true;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto next = [&](const char* flag) -> const char* {
            if (i + 1 >= argc) { fprintf(stderr, "tiny_renderer: %s needs a value\n", flag); exit(2); }   // reference: index panic
            return argv[++i];
        };
        if (a == "-p") asset_path = next("-p");
        else if (a == "-s") pipeline = next("-s");
        else if (a == "--out") out = next("--out");
        else if (a == "--width") width = (uint32_t)atoi(next("--width"));
        else if (a == "--height") height = (uint32_t)atoi(next("--height"));
        else if (a == "--camera-angle") camera_angle = strtof(next("--camera-angle"), nullptr);
        else if (a == "--light-angle") light_angle = strtof(next("--light-angle"), nullptr);
        else if (a == "--frames") frames = (uint32_t)atoi(next("--frames"));
        else if (a == "--camera-step") camera_step = strtof(next("--camera-step"), nullptr);
        else if (a == "--device") device = atoi(next("--device"));
        else if (a == "--dump-z") dump_z = next("--dump-z");
        else if (a == "--dump-shadow") dump_shadow = next("--dump-shadow");
        else if (a == "--replay") replay = next("--replay");
        else if (a == "--dt") fixed_dt = strtof(next("--dt"), nullptr);
        else if (a == "--views") n_views = (uint32_t)atoi(next("--views"));
        else if (a == "--lights") n_lights = (uint32_t)atoi(next("--lights"));
        else if (a == "--seed") seed = strtoull(next("--seed"), nullptr, 0);
        else if (a == "--gpus") n_gpus = (uint32_t)atoi(next("--gpus"));
        else if (a == "--ring") ring = (uint32_t)atoi(next("--ring"));
        else if (a == "--dump-visibility") dump_visibility = next("--dump-visibility");
        else if (a == "--quiet") print_fps = false;
        else if (a == "--help" || a == "-h") {
            printf("usage: tiny_renderer [-p asset_dir] [-s default|phong|normal_map|true_normal|specular|darboux|shadow|occlusion]\n"
                   "       [--out file.tga] [--width W] [--height H] [--camera-angle rad] [--light-angle rad]\n"
                   "       [--frames N --camera-step rad] [--dump-z file.tga] [--dump-shadow file.tga] [--device N] [--quiet]\n"
                   "       [--replay keys.txt [--dt seconds]]   one line per frame: keys pressed during that frame (a d = camera,\n"
                   "                                            q e = light, x = Escape released); the app loop of the reference, headless\n"
                   "       [--views N]    turntable batch (BASELINE config 5 shape): N views, camera angle --camera-angle + 2 pi k / N, light\n"
                   "                      --light-angle; prints the per-view checksums, --out receives the LAST view\n"
                   "       [--lights N [--seed S]]   many-light ambient occlusion (config 4; needs -s occlusion): N seeded light directions,\n"
                   "                      --out receives round(sum of the gray frames / N) as a gray image\n"
                   "       [--gpus G]     shard --views / --lights over devices 0..G-1 of this box (one process; tr_multi_*)\n"
                   "       [--ring R]     frame slots per scene for the batch modes (default 8 when batching)\n"
                   "       [--dump-visibility file]   winning polygon index per pixel, u32 little-endian, row 0 = bottom (0xFFFFFFFF = none)\n");
            return 0;
        }
        // unknown arguments are ignored, like the reference (main.rs:24)
    }
    printf("loading model from: %s/model.obj\n", asset_path.c_str());
    printf("cooking up a scene with '%s' shader pipeline\n", pipeline.c_str());
    std::vector<uint8_t> frame((size_t)width * height * 3);
    const float at[3] = {0, 0, 0}, up[3] = {0, 1, 0};
    if (n_views || n_lights) {
        // Batch modes (BASELINE configs 5 and 4) through the native multi-GPU driver: one scene per device, views /
        // lights in contiguous blocks per device, no Python and no NCCL (include/tinyr.h: tr_multi_*).
        if (n_gpus == 0 || n_gpus > 16) { fprintf(stderr, "tiny_renderer: --gpus 1..16\n"); return 2; }
        std::vector<int32_t> devices(n_gpus);
        for (uint32_t g = 0; g < n_gpus; ++g) devices[g] = device + (int32_t)g;
        tr_multi* multi = nullptr;
        if (tr_multi_create_from_dir(&multi, asset_path.c_str(), pipeline.c_str(), width, height, devices.data(), n_gpus, ring ? ring : 8)) return die("scenes");
        auto t0 = std::chrono::steady_clock::now();
        if (n_lights) {
            tr_mesh mesh{};
            if (tr_load_obj((asset_path + "/model.obj").c_str(), &mesh)) return die("model");
            std::vector<float> lights((size_t)n_lights * 3);
            uint64_t drawn = 0;
            if (tr_ao_generate_lights(mesh.positions, mesh.n_positions, at, up, seed, n_lights, 0.97f, lights.data(), &drawn)) return die("lights");
            tr_mesh_free(&mesh);
            const float from[3] = {sinf(camera_angle), 0.0f, cosf(camera_angle)};
            if (tr_multi_set_camera(multi, from, at, up)) return die("set_camera");
            std::vector<uint32_t> accum((size_t)width * height);
            if (tr_multi_render_ao(multi, lights.data(), n_lights, accum.data())) return die("render_ao");
            double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            // final image = accum / n, integer round-half-up; accum is bottom-up, the TGA writer wants top-down
            uint64_t checksum = 0;
            for (size_t i = 0; i < accum.size(); ++i) checksum += (uint64_t)accum[i] * (2ull * i + 1ull);
            for (uint32_t y = 0; y < height; ++y)
                for (uint32_t x = 0; x < width; ++x) {
                    const uint32_t g = (uint32_t)(((uint64_t)accum[(size_t)(height - 1 - y) * width + x] * 2 + n_lights) / (2ull * n_lights));
                    uint8_t* p = &frame[((size_t)y * width + x) * 3];
                    p[0] = p[1] = p[2] = (uint8_t)(g > 255 ? 255 : g);
                }
            printf("ao: %u lights (%llu drawn, seed 0x%llx) on %u device(s) in %.3f s (%.1f passes/s), checksum %llu\n", n_lights,
                   (unsigned long long)drawn, (unsigned long long)seed, n_gpus, secs, n_lights / secs, (unsigned long long)checksum);
        } else {
            std::vector<tr_view> views(n_views);
            for (uint32_t k = 0; k < n_views; ++k) {
                const float ca = camera_angle + 2.0f * 3.14159274101257324f * (float)k / (float)n_views;
                const tr_view v = {{sinf(ca), 0.0f, cosf(ca)}, {0, 0, 0}, {0, 1, 0}, {sinf(light_angle), 0.0f, cosf(light_angle)}};
                views[k] = v;
            }
            std::vector<uint64_t> sums(n_views);
            if (tr_multi_render_views(multi, views.data(), n_views, sums.data(), nullptr)) return die("render_views");
            double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            uint64_t all = 0;
            for (uint32_t k = 0; k < n_views; ++k) { if (print_fps) printf("view %u checksum %llu\n", k, (unsigned long long)sums[k]); all += sums[k]; }
            printf("views: %u on %u device(s) in %.3f s (%.1f frames/s, %.1f Mpixels/s), checksum of checksums %llu\n", n_views, n_gpus, secs,
                   n_views / secs, (double)n_views * width * height / secs / 1e6, (unsigned long long)all);
            // the device that rendered the last view still holds it
            if (tr_scene_get_frame_buffer(tr_multi_scene(multi, n_gpus - 1), frame.data())) return die("get_frame_buffer");
        }
        if (!out.empty() && tr_write_tga(out.c_str(), frame.data(), width, height, 1)) return die("write");
        if (!dump_visibility.empty()) {
            std::vector<uint32_t> vis((size_t)width * height);
            if (tr_scene_read_visibility(tr_multi_scene(multi, n_lights ? 0 : n_gpus - 1), vis.data())) return die("dump-visibility");
            FILE* fp = fopen(dump_visibility.c_str(), "wb");
            if (!fp || fwrite(vis.data(), 4, vis.size(), fp) != vis.size()) { fprintf(stderr, "tiny_renderer: cannot write %s\n", dump_visibility.c_str()); return 1; }
            fclose(fp);
        }
        tr_multi_destroy(multi);
        printf("wrote %s (%ux%u, %s)\n", out.c_str(), width, height, pipeline.c_str());
        return 0;
    }
    tr_scene* scene = nullptr;
    if (tr_scene_create_from_dir(&scene, asset_path.c_str(), pipeline.c_str(), width, height, device, 1)) return die("scene");
    auto t0 = std::chrono::steady_clock::now();
    if (!replay.empty()) {
        // Scripted interaction replay: the reference's app loop (app.rs:166-247) with the window's key events read from a
        // file.  Angles integrate SPEED * previous frame time (app.rs:173-199, CAMERA_SPEED = LIGHT_SOURCE_SPEED = 3.0,
        // app.rs:12-13); keys a/d move the camera, q/e the light, Escape (x) ends the loop at the start of the next frame
        // (app.rs:63-77, 224-227).  --dt fixes the frame time (deterministic replays); otherwise it is measured.
        FILE* fp = fopen(replay.c_str(), "r");
        if (!fp) { fprintf(stderr, "tiny_renderer: cannot open %s\n", replay.c_str()); return 1; }
        const float kCameraSpeed = 3.0f, kLightSpeed = 3.0f;
        bool cam_l = false, cam_r = false, light_l = false, light_r = false, exit_app = false;
        float frame_time = 0.0f;
        uint32_t shown = 0, fps_counter = 0;
        auto fps_begin = std::chrono::steady_clock::now();
        char line[256];
        while (!exit_app) {
            auto frame_begin = std::chrono::steady_clock::now();
            if (tr_scene_clear(scene)) return die("clear");
            if (cam_r) camera_angle += kCameraSpeed * frame_time;
            if (cam_l) camera_angle -= kCameraSpeed * frame_time;
            if (light_r) light_angle += kLightSpeed * frame_time;
            if (light_l) light_angle -= kLightSpeed * frame_time;
            const float from[3] = {sinf(camera_angle), 0.0f, cosf(camera_angle)};
            const float light[3] = {sinf(light_angle), 0.0f, cosf(light_angle)};
            if (tr_scene_set_light_direction(scene, light)) return die("set_light_direction");
            if (tr_scene_set_camera(scene, from, at, up)) return die("set_camera");
            if (tr_scene_render(scene)) return die("render");
            if (tr_scene_get_frame_buffer(scene, frame.data())) return die("get_frame_buffer");   // window.set_image in the reference
            ++shown;
            cam_l = cam_r = light_l = light_r = false;                                             // frame_action_buffer.reset()
            if (!fgets(line, sizeof line, fp)) exit_app = true;                                    // end of script = window closed
            else for (const char* c = line; *c; ++c) {
                if (*c == 'a') cam_l = true; else if (*c == 'd') cam_r = true;
                else if (*c == 'q') light_l = true; else if (*c == 'e') light_r = true;
                else if (*c == 'x') exit_app = true;
            }
            if (print_fps) {
                ++fps_counter;
                if (std::chrono::duration<float>(std::chrono::steady_clock::now() - fps_begin).count() > 1.0f) {
                    printf("FPS --- %u\n", fps_counter);                                          // app.rs:238
                    fps_begin = std::chrono::steady_clock::now();
                    fps_counter = 0;
                }
            }
            frame_time = fixed_dt >= 0.0f ? fixed_dt : std::chrono::duration<float>(std::chrono::steady_clock::now() - frame_begin).count();
        }
        fclose(fp);
        double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        printf("replayed %u frames in %.3f s (%.1f frames/s), camera angle %.9g, light angle %.9g\n", shown, secs, shown / secs,
               (double)camera_angle, (double)light_angle);
        frames = 0;
    }
    for (uint32_t f = 0; f < frames; ++f) {
        float ca = camera_angle + camera_step * (float)f;
        const float from[3] = {sinf(ca), 0.0f, cosf(ca)};
        const float light[3] = {sinf(light_angle), 0.0f, cosf(light_angle)};
        if (tr_scene_clear(scene)) return die("clear");                                   // app.rs:170
        if (tr_scene_set_light_direction(scene, light)) return die("set_light_direction"); // app.rs:208
        if (tr_scene_set_camera(scene, from, at, up)) return die("set_camera");            // app.rs:209
        if (tr_scene_render(scene)) return die("render");                                  // app.rs:210
        if (tr_scene_get_frame_buffer(scene, frame.data())) return die("get_frame_buffer"); // app.rs:213
    }
    double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (print_fps && frames > 1) printf("FPS --- %.1f\n", frames / secs);                   // app.rs:238
    if (!out.empty() && tr_write_tga(out.c_str(), frame.data(), width, height, 1)) return die("write");
    if (!dump_z.empty()) {
        if (tr_scene_get_z_buffer(scene, frame.data()) || tr_write_tga(dump_z.c_str(), frame.data(), width, height, 1)) return die("dump-z");
    }
    if (!dump_shadow.empty()) {
        if (tr_scene_get_shadow_buffer(scene, frame.data()) || tr_write_tga(dump_shadow.c_str(), frame.data(), width, height, 1)) return die("dump-shadow");
    }
    if (!dump_visibility.empty()) {
        std::vector<uint32_t> vis((size_t)width * height);
        if (tr_scene_read_visibility(scene, vis.data())) return die("dump-visibility");
        FILE* fp = fopen(dump_visibility.c_str(), "wb");
        if (!fp || fwrite(vis.data(), 4, vis.size(), fp) != vis.size()) { fprintf(stderr, "tiny_renderer: cannot write %s\n", dump_visibility.c_str()); return 1; }
        fclose(fp);
    }
    uint64_t diag[8];
    if (tr_scene_diagnostics(scene, diag) == 0 && (diag[0] | diag[2] | diag[4] | diag[5]))
        fprintf(stderr, "tiny_renderer: warning: the reference would have panicked on this input (shadow oob %llu, texel oob %llu, w==0 %llu, singular basis %llu)\n",
                (unsigned long long)diag[0], (unsigned long long)diag[2], (unsigned long long)diag[4], (unsigned long long)diag[5]);
    tr_scene_destroy(scene);
    printf("wrote %s (%ux%u, %s)\n", out.c_str(), width, height, pipeline.c_str());
    return 0;
}

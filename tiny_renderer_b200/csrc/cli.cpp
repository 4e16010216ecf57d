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
    std::string asset_path = "assets/diablo", pipeline = "default", out = "out.tga", dump_z, dump_shadow, replay;
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
        else if (a == "--quiet") print_fps = false;
        else if (a == "--help" || a == "-h") {
            printf("usage: tiny_renderer [-p asset_dir] [-s default|phong|normal_map|true_normal|specular|darboux|shadow|occlusion]\n"
                   "       [--out file.tga] [--width W] [--height H] [--camera-angle rad] [--light-angle rad]\n"
                   "       [--frames N --camera-step rad] [--dump-z file.tga] [--dump-shadow file.tga] [--device N] [--quiet]\n"
                   "       [--replay keys.txt [--dt seconds]]   one line per frame: keys pressed during that frame (a d = camera,\n"
                   "                                            q e = light, x = Escape released); the app loop of the reference, headless\n");
            return 0;
        }
        // unknown arguments are ignored, like the reference (main.rs:24)
    }
    printf("loading model from: %s/model.obj\n", asset_path.c_str());
    printf("cooking up a scene with '%s' shader pipeline\n", pipeline.c_str());
    tr_scene* scene = nullptr;
    if (tr_scene_create_from_dir(&scene, asset_path.c_str(), pipeline.c_str(), width, height, device, 1)) return die("scene");
    std::vector<uint8_t> frame((size_t)width * height * 3);
    const float at[3] = {0, 0, 0}, up[3] = {0, 1, 0};
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
    uint64_t diag[8];
    if (tr_scene_diagnostics(scene, diag) == 0 && (diag[0] | diag[2] | diag[4] | diag[5]))
        fprintf(stderr, "tiny_renderer: warning: the reference would have panicked on this input (shadow oob %llu, texel oob %llu, w==0 %llu, singular basis %llu)\n",
                (unsigned long long)diag[0], (unsigned long long)diag[2], (unsigned long long)diag[4], (unsigned long long)diag[5]);
    tr_scene_destroy(scene);
    printf("wrote %s (%ux%u, %s)\n", out.c_str(), width, height, pipeline.c_str());
    return 0;
}

// input_replay.h — the per-frame input handling of the reference's app loop (VoxelApp/main.cu:70-161) as a plain function
// driven by a recorded script instead of SDL events, plus its frame-time average (main.cu:173-183).  Host-only C++ (float3 /
// float2 come from CUDA's vector_types.h, which is plain C); used by examples/vxview_main.cu and tested on the CPU.
//
// One script line per frame:   <keys> <mouse_x> <mouse_y> <click> <wheel>
//   keys   letters of the keys held this frame out of W A S D Q E, '^' for left shift, '-' for none   (main.cu:107-143)
//   mouse  pointer position in window pixels (SDL_GetMouseState, main.cu:150)
//   click  1 while the left button is down (main.cu:78-91)
//   wheel  +n / -n wheel notches received this frame (each is one SDL_MOUSEWHEEL event, main.cu:93-105)
#pragma once
#include <vector_types.h>
#include <cstdio>
#include <cstring>
#include <string>

namespace replay {

struct InputFrame {
    bool w = false, a = false, s = false, d = false, q = false, e = false, shift = false, click = false;
    int mouse_x = 0, mouse_y = 0, wheel = 0;
};

struct CameraRig { // the locals of main() that the callback captures (main.cu:51-55, :64, :147)
    float3 pos{256, 256, 256};
    float3 up{0, 1, 0}, right{1, 0, 0}, forward{0, 0, 1}, eular{0, 0, 0};
    float2 ortho{10, 10};
    int last_x = 0, last_y = 0;
};

inline bool parse_line(const char* line, InputFrame& f) {
    char keys[32];
    int x = 0, y = 0, click = 0, wheel = 0;
    if (sscanf(line, "%31s %d %d %d %d", keys, &x, &y, &click, &wheel) != 5) return false;
    f = InputFrame();
    for (const char* k = keys; *k; ++k) {
        switch (*k) {
        case 'W': case 'w': f.w = true; break;
        case 'A': case 'a': f.a = true; break;
        case 'S': case 's': f.s = true; break;
        case 'D': case 'd': f.d = true; break;
        case 'Q': case 'q': f.q = true; break;
        case 'E': case 'e': f.e = true; break;
        case '^': f.shift = true; break;
        case '-': break;
        default: return false;
        }
    }
    f.mouse_x = x; f.mouse_y = y; f.click = click != 0; f.wheel = wheel;
    return true;
}

// helper_math.h's float3 * float and += / -= (component-wise binary32 products and sums)
inline void move(float3& p, const float3& dir, float speed, float sign) {
    p.x += sign * (dir.x * speed); p.y += sign * (dir.y * speed); p.z += sign * (dir.z * speed);
}

// What one invocation of the render callback does to the camera before GetDirections / RenderScreen (main.cu:72-158).
// The basis vectors are those of the PREVIOUS frame, as in the reference (GetDirections runs after the moves).
// Returns true when the ortho window changed (the caller passes rig.ortho to SetOrthoWindowSize either way, :105, :121).
inline bool apply_input(CameraRig& rig, const InputFrame& in) {
    bool ortho_changed = false;
    for (int n = in.wheel; n != 0; n += (n > 0 ? -1 : 1)) { // one event per notch
        const float step = n > 0 ? -10.0f : 10.0f;          // scroll up shrinks the window
        rig.ortho.x += step; rig.ortho.y += step;
        ortho_changed = true;
    }
    float cam_speed = 0.2; // `float cam_speed = 0.2;` — the double literal narrowed
    if (in.shift) cam_speed *= 10;
    if (in.w) move(rig.pos, rig.forward, cam_speed, 1.0f);
    if (in.s) move(rig.pos, rig.forward, cam_speed, -1.0f);
    if (in.a) move(rig.pos, rig.right, cam_speed, -1.0f);
    if (in.d) move(rig.pos, rig.right, cam_speed, 1.0f);
    if (in.q) move(rig.pos, rig.up, cam_speed, -1.0f);
    if (in.e) move(rig.pos, rig.up, cam_speed, 1.0f);
    if (in.click) { // mouse look: radians per pixel 0.004, pitch from dy, yaw from dx
        const int dx = in.mouse_x - rig.last_x, dy = in.mouse_y - rig.last_y;
        rig.eular.x += dy * 0.004f;
        rig.eular.y += dx * 0.004f;
    }
    rig.last_x = in.mouse_x; rig.last_y = in.mouse_y;
    return ortho_changed;
}

// main.cu:173-183: exponential average of the frame time (ms), weight 1/100, starting from 0
struct FrameClock {
    double avg_ms = 0.0;
    double add(double ms) { avg_ms = avg_ms * (1 - (1 / 100.0)) + ms * (1 / 100.0); return avg_ms; }
    double avg_fps() const { return 1000.0 / avg_ms; }
};

// BGRA8 frame (PixelData as the reference's kernels write it, Renderer.cu:72-87) -> binary PPM
inline bool write_ppm(const std::string& path, const unsigned char* bgra, unsigned width, unsigned height) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) return false;
    fprintf(f, "P6\n%u %u\n255\n", width, height);
    std::string row(width * 3, '\0');
    for (unsigned y = 0; y < height; ++y) {
        const unsigned char* p = bgra + (size_t)y * width * 4;
        for (unsigned x = 0; x < width; ++x) { row[3 * x] = (char)p[4 * x + 2]; row[3 * x + 1] = (char)p[4 * x + 1]; row[3 * x + 2] = (char)p[4 * x]; }
        fwrite(row.data(), 1, row.size(), f);
    }
    return fclose(f) == 0;
}

} // namespace replay

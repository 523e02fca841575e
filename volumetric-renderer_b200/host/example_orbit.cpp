// example_orbit.cpp -- the host loop of the application reduced to this path, in C++ over the
// mirror classes: create the pass, load / generate a volume, push a transfer function, then per
// frame rotate the camera (Camera::rotate as src/ui/main_window.cpp:289 does), record, and read
// the image back.  Headless: the frame is written as a binary PPM instead of being shown.
//
//   example_orbit out.ppm [W H n frames [file.nrrd]]
//
// Build: g++ -std=c++20 example_orbit.cpp -L.. -lvr_b200 -Wl,-rpath,'$ORIGIN/..' -o example_orbit
#include <cstdio>
#include <cstdlib>
#include <fstream>

#include "offscreen_pass.hpp"

int main(int argc, char **argv)
{
    const char *out = argc > 1 ? argv[1] : "orbit.ppm";
    const uint32_t W = argc > 2 ? std::atoi(argv[2]) : 512, H = argc > 3 ? std::atoi(argv[3]) : 512;
    const uint32_t n = argc > 4 ? std::atoi(argv[4]) : 128;
    const int frames = argc > 5 ? std::atoi(argv[5]) : 8;
    try {
        VolB200::Rendering::OffscreenPass pass(0, W, H);
        if (argc > 6) pass.import_nrrd(argv[6]);
        else if (vr_synth_volume(pass.handle(), VR_SYNTH_FBM, 0x5EED0002u, VR_U16, n, n, n) != VR_OK)
            throw std::runtime_error(vr_last_error(pass.handle()));
        VolB200::UI::Components::Gradient g;  // TF-B of SURVEY.md 8(d)
        g.color_markers = {{0.0f, {0.0f, 0.0f, 0.5f}}, {0.5f, {1.0f, 0.5f, 0.0f}}, {1.0f, {1.0f, 1.0f, 1.0f}}};
        g.alpha_markers = {{0.0f, 0.0f}, {0.3f, 0.0f}, {1.0f, 0.08f}};
        pass.transfer_function_changed(g.discretize(256));
        pass.slicing_changed({0.0f, 0.0f, 0.0f}, {1.0f, 1.0f, 1.0f});
        VolB200::Scene::Camera camera;
        std::vector<uint8_t> img;
        unsigned long long sum = 0;
        for (int k = 0; k < frames; k++) {
            pass.record(nullptr, 0, &camera);
            img = pass.download();
            for (uint8_t b : img) sum += b;
            camera.rotate({4.0f, 0.0f});
        }
        std::ofstream f(out, std::ios::binary);
        f << "P6\n" << W << " " << H << "\n255\n";
        for (size_t i = 0; i < (size_t)W * H; i++) f.write(reinterpret_cast<const char *>(&img[4 * i]), 3);
        std::printf("frames=%d checksum=%llu last_frame=%s\n", frames, sum, out);
    } catch (const std::exception &e) {
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    return 0;
}

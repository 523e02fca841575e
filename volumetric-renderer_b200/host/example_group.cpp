// example_group.cpp -- example_orbit.cpp on several GPUs of one process: the same host loop over
// VolB200::Rendering::PassGroup (vr_create_group), sort-first or sort-last, next to a single pass on
// the first device; prints both checksums.  Devices may repeat (logical ranks on one GPU).
//
//   example_group sort_first|sort_last W H n frames dev [dev ...]
//
// Build: g++ -std=c++20 example_group.cpp -L.. -lvr_b200 -Wl,-rpath,'$ORIGIN/..' -o example_group
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "offscreen_pass.hpp"

int main(int argc, char **argv)
{
    if (argc < 7) {
        std::fprintf(stderr, "usage: example_group sort_first|sort_last W H n frames dev [dev ...]\n");
        return 2;
    }
    const vr_partition part = std::strcmp(argv[1], "sort_last") == 0 ? VR_PARTITION_SORT_LAST : VR_PARTITION_SORT_FIRST;
    const uint32_t W = std::atoi(argv[2]), H = std::atoi(argv[3]), n = std::atoi(argv[4]);
    const int frames = std::atoi(argv[5]);
    std::vector<int> devs;
    for (int i = 6; i < argc; i++) devs.push_back(std::atoi(argv[i]));
    try {
        using namespace VolB200;
        UI::Components::Gradient g;  // TF-B of SURVEY.md 8(d)
        g.color_markers = {{0.0f, {0.0f, 0.0f, 0.5f}}, {0.5f, {1.0f, 0.5f, 0.0f}}, {1.0f, {1.0f, 1.0f, 1.0f}}};
        g.alpha_markers = {{0.0f, 0.0f}, {0.3f, 0.0f}, {1.0f, 0.08f}};
        const std::vector<uint32_t> lut = g.discretize(256);
        Rendering::PassGroup group(devs, part, W, H, 2);
        group.synth_volume(VR_SYNTH_FBM, 0x5EED0002u, VR_U16, n);
        group.transfer_function_changed(lut);
        Rendering::OffscreenPass single(devs[0], W, H);
        if (vr_synth_volume(single.handle(), VR_SYNTH_FBM, 0x5EED0002u, VR_U16, n, n, n) != VR_OK)
            throw std::runtime_error(vr_last_error(single.handle()));
        single.transfer_function_changed(lut);
        if (part == VR_PARTITION_SORT_LAST) {
            // the merged frame is compared with one GPU's: both without early termination, 3-D array everywhere
            vr_options o;
            vr_get_options(single.handle(), &o);
            o.early_term = 0;
            vr_set_options(single.handle(), &o);
            group.set_options(o);
        }
        Scene::Camera camera;
        unsigned long long sum_group = 0, sum_single = 0;
        long worst = 0;
        for (int k = 0; k < frames; k++) {
            group.record(&camera);
            single.record(nullptr, 0, &camera);
            const std::vector<uint8_t> a = group.download(), b = single.download();
            for (size_t i = 0; i < a.size(); i++) {
                sum_group += a[i], sum_single += b[i];
                const long d = std::labs((long)a[i] - (long)b[i]);
                worst = d > worst ? d : worst;
            }
            camera.rotate({4.0f, 0.0f});
        }
        std::printf("ranks=%u frames=%d checksum_group=%llu checksum_single=%llu max_abs_diff=%ld\n", group.size(), frames, sum_group,
                    sum_single, worst);
    } catch (const std::exception &e) {
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    return 0;
}

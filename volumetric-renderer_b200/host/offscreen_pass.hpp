// offscreen_pass.hpp -- C++ host side above the C ABI, mirroring the reference's own classes for
// this path so that the application's call sites compile against it unchanged in meaning:
//
//   Vol::Rendering::OffscreenPass   src/rendering/offscreen_pass.h:27-148  (public members :39-54)
//   Vol::Scene::Camera              src/scene/camera.h:9-23
//   Vol::Data::Dataset              src/data/dataset.h:9-13
//   Vol::UI::Components::Gradient   src/ui/components/gradient.h:11-36  (discretize only)
//
// Header-only; links against libvr_b200.so (include/vr.h).  Like the reference, failures throw
// std::runtime_error (e.g. offscreen_pass.cpp:309-313); the message is vr_last_error().
// No glm dependency: vectors are std::array<float, N>, matrices column-major float[16] (glm layout).
#pragma once

#include <array>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../../include/vr.h"

namespace VolB200
{
using vec2 = std::array<float, 2>;
using vec3 = std::array<float, 3>;
using mat4 = std::array<float, 16>;

namespace Data
{
// dataset.h:9-13.  `data` as the reference holds it (floats); native u8/u16 payloads can be
// handed over with OffscreenPass::volume_dataset_changed(ptr, dtype, ...) instead.
struct Dataset {
    std::array<uint32_t, 3> dimensions;
    float min, max;
    std::vector<float> data;
};
}  // namespace Data

namespace Scene
{
class Camera
{
  public:
    explicit Camera() : h(vr_camera_new())
    {
        if (!h) throw std::runtime_error("vr_camera_new failed");
    }
    ~Camera() { vr_camera_free(h); }
    Camera(const Camera &) = delete;
    Camera &operator=(const Camera &) = delete;

    void rotate(const vec2 &delta) { vr_camera_rotate(h, delta[0], delta[1]); }  // camera.cpp:15-29
    void zoom(float delta) { vr_camera_zoom(h, delta); }                          // camera.cpp:31-34
    vec3 get_position() const                                                      // camera.cpp:36-40
    {
        vec3 p;
        vr_camera_get_position(h, p.data());
        return p;
    }
    mat4 get_view() const  // camera.cpp:42-48
    {
        mat4 m;
        vr_camera_get_view(h, m.data());
        return m;
    }
    const vr_camera *handle() const { return h; }

  private:
    vr_camera *h;
};
}  // namespace Scene

namespace UI::Components
{
struct Gradient {
    template <typename T>
    struct Marker {
        float location;
        T value;
    };
    std::vector<Marker<vec3>> color_markers{{0.0f, {0.0f, 0.0f, 0.0f}}, {1.0f, {1.0f, 1.0f, 1.0f}}};  // gradient.cpp:64-70
    std::vector<Marker<float>> alpha_markers{{0.0f, 1.0f}, {1.0f, 1.0f}};

    std::vector<uint32_t> discretize(size_t count) const  // gradient.cpp:90-108
    {
        std::vector<float> cm, am;
        for (const auto &m : color_markers) cm.insert(cm.end(), {m.location, m.value[0], m.value[1], m.value[2]});
        for (const auto &m : alpha_markers) am.insert(am.end(), {m.location, m.value});
        std::vector<uint32_t> out(count);
        if (vr_gradient_discretize(cm.data(), (uint32_t)color_markers.size(), am.data(), (uint32_t)alpha_markers.size(),
                                   (uint32_t)count, out.data()) != VR_OK)
            throw std::runtime_error("vr_gradient_discretize: bad markers");
        return out;
    }
};
}  // namespace UI::Components

namespace Rendering
{
class OffscreenPass
{
  public:
    // OffscreenPass(VulkanContext*, width, height): the CUDA device index stands where the
    // Vulkan context stood (offscreen_pass.cpp:112-134).
    explicit OffscreenPass(int device, uint32_t width, uint32_t height)
    {
        if (vr_create(device, width, height, &ctx) != VR_OK) throw std::runtime_error(vr_last_error(nullptr));
    }
    ~OffscreenPass() { vr_destroy(ctx); }
    OffscreenPass(const OffscreenPass &) = delete;
    OffscreenPass &operator=(const OffscreenPass &) = delete;

    // record(command_buffer, frame_index): the stream stands where the command buffer stood.
    // The reference pulls the camera inside record() (update_uniform_buffer, :1152-1171);
    // pass it here to keep that behaviour.
    void record(void *cuda_stream, uint32_t frame_index, const Scene::Camera *camera = nullptr)
    {
        if (camera) check(vr_set_camera_orbit(ctx, camera->handle()));
        check(vr_render_frame(ctx, frame_index, cuda_stream));
    }
    void framebuffer_size_changed(uint32_t width, uint32_t height) { check(vr_resize(ctx, width, height)); }
    void volume_dataset_changed(Data::Dataset &dataset)
    {
        check(vr_set_volume(ctx, dataset.data.data(), VR_F32, dataset.dimensions[0], dataset.dimensions[1],
                            dataset.dimensions[2], &dataset.min, &dataset.max));
    }
    // native-type overload: what the loader fast path hands over (min/max computed on the device)
    void volume_dataset_changed(const void *voxels, vr_dtype dtype, uint32_t nx, uint32_t ny, uint32_t nz)
    {
        check(vr_set_volume(ctx, voxels, dtype, nx, ny, nz, nullptr, nullptr));
    }
    void slicing_changed(const vec3 &min, const vec3 &max) { check(vr_set_slicing(ctx, min.data(), max.data())); }
    void transfer_function_changed(const std::vector<uint32_t> &data)
    {
        check(vr_set_transfer_function(ctx, data.data(), (uint32_t)data.size()));
    }
    // the gradient editor's edit (main_window.cpp:250-257) without the host-side discretize: built on the device
    void transfer_function_changed(const UI::Components::Gradient &gradient, uint32_t count = 256)
    {
        std::vector<float> cm, am;
        for (const auto &m : gradient.color_markers) cm.insert(cm.end(), {m.location, m.value[0], m.value[1], m.value[2]});
        for (const auto &m : gradient.alpha_markers) am.insert(am.end(), {m.location, m.value});
        check(vr_set_gradient(ctx, cm.data(), (uint32_t)gradient.color_markers.size(), am.data(), (uint32_t)gradient.alpha_markers.size(), count));
    }
    // honour a file's voxel spacings (the reference ignores them): nullptr = the unit cube
    void set_volume_spacing(const vec3 *spacing) { check(vr_set_volume_spacing(ctx, spacing ? spacing->data() : nullptr)); }
    // optional quality mode, no reference counterpart: classify segments between samples from a 2-D table (0 = off)
    void set_preintegration(uint32_t substeps) { check(vr_set_preintegration(ctx, substeps)); }
    // get_image_view() / get_sampler(): the colour attachment as a device pointer + pitch
    std::pair<void *, size_t> get_image_view(uint32_t frame_index = 0) const
    {
        void *p = nullptr;
        size_t pitch = 0;
        check(vr_get_image_frame(ctx, frame_index, &p, &pitch));
        return {p, pitch};
    }

    // Importer::import(FileFormat::Nrrd) without the dialog (importer.cpp:36-46)
    void import_nrrd(const std::string &path) { check(vr_load_nrrd(ctx, path.c_str())); }
    // Importer::import(FileFormat::CSV) without the dialog (importer.cpp:30-46): one file per z slice
    void import_csv(const std::vector<std::string> &paths)
    {
        std::vector<const char *> p;
        for (const std::string &s : paths) p.push_back(s.c_str());
        check(vr_load_csv(ctx, p.data(), (uint32_t)p.size()));
    }
    void set_uniforms(const mat4 &view, const mat4 &proj, const vec3 &camera_position)
    {
        check(vr_set_camera(ctx, view.data(), proj.data(), camera_position.data()));
    }
    std::vector<uint8_t> download()
    {
        uint32_t w = 0, h = 0;
        vr_get_size(ctx, &w, &h);
        std::vector<uint8_t> img((size_t)w * h * 4);
        check(vr_download(ctx, img.data()));
        return img;
    }
    // The viewport's sampling of the image (LINEAR, CLAMP_TO_EDGE; offscreen_pass.cpp:361-375, drawn by
    // main_window.cpp:157-165): the frame resampled into a device RGBA8 buffer of another size.
    void present(void *dst_device_rgba8, uint32_t width, uint32_t height, size_t pitch_bytes = 0, void *cuda_stream = nullptr,
                 uint32_t frame_index = 0)
    {
        check(vr_present_frame(ctx, frame_index, dst_device_rgba8, width, height, pitch_bytes, cuda_stream));
    }
    // Bring a host RGBA8 buffer up to date with a frame in flight, copying only what can have changed since the
    // last delivery into that buffer (outside the cube's screen box a frame is the clear colour, :170-173).
    size_t download_dirty_async(void *host_rgba8, void *cuda_stream = nullptr, uint32_t frame_index = 0)
    {
        size_t bytes = 0;
        check(vr_download_frame_dirty_async(ctx, frame_index, host_rgba8, cuda_stream, &bytes));
        return bytes;
    }
    void write_image(const std::string &path) { check(vr_write_image(ctx, path.c_str())); }
    void wait_idle() { check(vr_sync(ctx)); }
    vr_ctx *handle() { return ctx; }

  private:
    void check(int rc) const
    {
        if (rc != VR_OK) throw std::runtime_error(vr_last_error(ctx));
    }
    vr_ctx *ctx = nullptr;
};

// One pass spread over several GPUs of one process (include/vr.h, vr_create_group).  It stands where the
// application constructs and records its single pass (src/rendering/vulkan_context.cpp:51, main_pass.cpp:91) and
// keeps OffscreenPass's member names; mutators reach every device, record() enqueues one frame on all of them and
// returns, the finished frame is the display rank's (rank 0) image.
class PassGroup
{
  public:
    PassGroup(const std::vector<int> &devices, vr_partition partition, uint32_t width, uint32_t height, uint32_t slots = 2) : w(width), h(height)
    {
        if (vr_create_group((uint32_t)devices.size(), devices.data(), partition, width, height, slots, &g) != VR_OK)
            throw std::runtime_error(vr_group_last_error(nullptr));
    }
    ~PassGroup() { vr_group_destroy(g); }
    PassGroup(const PassGroup &) = delete;
    PassGroup &operator=(const PassGroup &) = delete;

    uint32_t size() const { return vr_group_size(g); }
    // record(command_buffer, frame_index): frames in flight are the group's own slots, taken in turn
    void record(const Scene::Camera *camera = nullptr)
    {
        if (camera) {
            mat4 view, proj;
            vr_camera_get_view(camera->handle(), view.data());
            vr_projection(w, h, proj.data());
            set_uniforms(view, proj, camera->get_position());
        }
        check(vr_group_render(g));
    }
    void set_uniforms(const mat4 &view, const mat4 &proj, const vec3 &camera_position)
    {
        check(vr_group_set_camera(g, view.data(), proj.data(), camera_position.data()));
    }
    void volume_dataset_changed(Data::Dataset &dataset)
    {
        check(vr_group_set_volume(g, dataset.data.data(), VR_F32, dataset.dimensions[0], dataset.dimensions[1], dataset.dimensions[2],
                                  &dataset.min, &dataset.max));
    }
    void volume_dataset_changed(const void *voxels, vr_dtype dtype, uint32_t nx, uint32_t ny, uint32_t nz)
    {
        check(vr_group_set_volume(g, voxels, dtype, nx, ny, nz, nullptr, nullptr));
    }
    void synth_volume(int kind, uint32_t seed, vr_dtype dtype, uint32_t n) { check(vr_group_synth_volume(g, kind, seed, dtype, n, n, n)); }
    void slicing_changed(const vec3 &min, const vec3 &max) { check(vr_group_set_slicing(g, min.data(), max.data())); }
    void transfer_function_changed(const std::vector<uint32_t> &data)
    {
        check(vr_group_set_transfer_function(g, data.data(), (uint32_t)data.size()));
    }
    void set_options(const vr_options &o) { check(vr_group_set_options(g, &o)); }
    // sort-last: how far the split planes may follow the camera (fraction of the edge; before the volume is set)
    void set_balance(float fraction) { check(vr_group_set_balance(g, fraction)); }
    // the frame recorded last, complete (replicas: of `rank`)
    std::vector<uint8_t> download(uint32_t rank = 0)
    {
        std::vector<uint8_t> img((size_t)w * h * 4);
        check(vr_group_download(g, rank, img.data()));
        return img;
    }
    void wait_idle() { check(vr_group_sync(g)); }
    vr_ctx *context(uint32_t rank) { return vr_group_context(g, rank); }
    vr_group *handle() { return g; }

  private:
    void check(int rc) const
    {
        if (rc != VR_OK) throw std::runtime_error(vr_group_last_error(g));
    }
    vr_group *g = nullptr;
    uint32_t w, h;
};
}  // namespace Rendering
}  // namespace VolB200

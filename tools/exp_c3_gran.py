"""tools/exp_c3_gran.py: config 3, how many pixels / 8x4 warp tiles / 16x8 CTA tiles of a frame draw anything, against
what the tile mask marks.  Says what a finer mask could still drop.  Run under gpurun."""
import sys, numpy as np
sys.path.insert(0, ".")
import volumetric_renderer_b200 as vr
import bench
op = vr.OffscreenPass(3840, 2160)
op.synth_volume(vr.SYNTH_BLOBS, 0x5EED0003, vr.U16, 1024)
op.transfer_function_changed(bench.product_lut(vr, "C"))
op.set_options(sampler=vr.SAMPLER_HW, empty_skip=1, early_term=1, early_term_eps=1/1024, write_raw=1)
print("occupancy", op.occupancy_info() if hasattr(op, "occupancy_info") else "")
for k in (0, 23, 45):
    cam = vr.Camera()
    for _ in range(k): cam.rotate((4.0, 0.0))
    op.set_camera(cam); op.record()
    raw = op.download_raw()
    drew = raw[..., 3] != 0
    bits, ign, th = op.tile_mask(0)
    H, W = drew.shape
    def blocks(bw, bh):
        return drew[:H // bh * bh, :W // bw * bw].reshape(H // bh, bh, W // bw, bw).any(axis=(1, 3))
    t16 = blocks(16, 8); t8 = blocks(8, 4); t4 = blocks(4, 4)
    marked_px = int(bits.sum()) * 16 * th
    print(f"frame {k}: drew px {int(drew.sum())}; marked tiles {int(bits.sum())} ({marked_px} px); tiles that drew 16x8 {int(t16.sum())} ({int(t16.sum())*128} px), "
          f"8x4 {int(t8.sum())} ({int(t8.sum())*32} px), 4x4 {int(t4.sum())} ({int(t4.sum())*16} px)")
    # warps (8x4) inside marked tiles
    up = np.repeat(np.repeat(bits, 2, axis=0), 2, axis=1)[:H // 4, :W // 8]
    print(f"          8x4 warp tiles inside marked CTA tiles {int(up.sum())}, of which drew {int((up & t8).sum())}")

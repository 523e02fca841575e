"""Shared helpers of the test-suite: parity metrics, transfer functions, cameras."""
import numpy as np

from oracle import pyoracle

# BASELINE.json north_star: <= 2/255 per RGBA channel at the 99.9th-percentile pixel, >= 45 dB PSNR
TOL_P999 = 2.0
TOL_PSNR = 45.0


def p999(a: np.ndarray, b: np.ndarray) -> float:
    """99.9th-percentile over pixels of the per-pixel max-channel absolute difference (0..255)."""
    d = np.abs(a.astype(np.int16) - b.astype(np.int16)).max(axis=-1).ravel()
    return float(np.percentile(d, 99.9))


def psnr(a: np.ndarray, b: np.ndarray) -> float:
    mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
    return float("inf") if mse == 0 else float(10.0 * np.log10(255.0 ** 2 / mse))


def assert_within_tolerance(img, ref_img, what=""):
    e, s = p999(img, ref_img), psnr(img, ref_img)
    assert e <= TOL_P999 and s >= TOL_PSNR, f"{what}: p99.9={e} (<= {TOL_P999}), PSNR={s:.2f} dB (>= {TOL_PSNR})"
    return e, s


def bits(a: np.ndarray) -> np.ndarray:
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


TF_0 = pyoracle.TF_0


def tf(oracle, name: str, count: int = 256) -> np.ndarray:
    if name == "0":
        return TF_0.copy()
    markers = {"A": pyoracle.TF_A_MARKERS, "B": pyoracle.TF_B_MARKERS, "C": pyoracle.TF_C_MARKERS}[name]
    return oracle.discretize(*markers, count=count)


def vr_gradient(vr, name: str):
    cm, am = {"A": pyoracle.TF_A_MARKERS, "B": pyoracle.TF_B_MARKERS, "C": pyoracle.TF_C_MARKERS}[name]
    return vr.Gradient([(m[0], tuple(m[1:])) for m in cm], [tuple(m) for m in am])


def twin_cameras(vr, oracle, moves=(), zoom=0.0):
    """The product camera and the oracle camera after the same mouse input."""
    a, b = vr.Camera(), oracle.camera()
    for m in moves:
        a.rotate(m)
        b.rotate(*m)
    if zoom:
        a.zoom(zoom)
        b.zoom(zoom)
    return a, b


def oracle_frame(oracle, cam, w, h, mn, mx, min_slice=(0, 0, 0), max_slice=(1, 1, 1)):
    return oracle.frame(w, h, cam, mn, mx, tuple(float(x) for x in min_slice), tuple(float(x) for x in max_slice))

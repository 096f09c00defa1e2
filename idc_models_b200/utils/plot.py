"""History logging — ``log(path, history, history_fine, num_devices)`` of the transfer scripts
(/root/reference/dist_model_tf_mobile.py:67-101): concatenates the phase-1 and
phase-2 curves, marks the start of fine-tuning, writes ``<path>/logs/plot_dev{N}.png`` and prints both raw
``history`` dicts.

matplotlib is not installed on this image, so the PNG is rendered by a tiny built-in rasteriser (poly-lines
into an RGB array, encoded with zlib — no external dependency); when matplotlib IS importable it is used
with the reference's exact figure layout. A ``.json`` with the raw numbers is always written next to it.
"""
from __future__ import annotations

import json
import os
import struct
import zlib

import numpy as np


def _png_bytes(img: np.ndarray) -> bytes:
    h, w, _ = img.shape
    raw = b"".join(b"\x00" + img[y].tobytes() for y in range(h))

    def chunk(tag, data):
        c = struct.pack(">I", len(data)) + tag + data
        return c + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)
    return (b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) +
            chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def _line(img, x0, y0, x1, y1, color):
    n = int(max(abs(x1 - x0), abs(y1 - y0), 1))
    xs = np.linspace(x0, x1, n + 1).round().astype(int)
    ys = np.linspace(y0, y1, n + 1).round().astype(int)
    ok = (xs >= 0) & (xs < img.shape[1]) & (ys >= 0) & (ys < img.shape[0])
    img[ys[ok], xs[ok]] = color


def _panel(img, top, height, series, ylim, marker_x, n_epochs):
    left, right = 60, img.shape[1] - 20
    bottom = top + height
    _line(img, left, top, left, bottom, (0, 0, 0))
    _line(img, left, bottom, right, bottom, (0, 0, 0))
    lo, hi = ylim
    span = max(hi - lo, 1e-12)

    def px(i, v):
        x = left + (right - left) * (i / max(n_epochs - 1, 1))
        y = bottom - (bottom - top) * ((min(max(v, lo), hi) - lo) / span)
        return x, y
    colors = [(31, 119, 180), (255, 127, 14)]
    for si, vals in enumerate(series):
        for i in range(len(vals) - 1):
            x0, y0 = px(i, vals[i]); x1, y1 = px(i + 1, vals[i + 1])
            _line(img, x0, y0, x1, y1, colors[si % 2])
    mx, _ = px(marker_x, lo)
    _line(img, mx, top, mx, bottom, (44, 160, 44))


def log(path: str, history, history_fine, num_devices: int, initial_epochs: int = 10, acc_ylim=(0.8, 1.0),
        loss_ylim=(0.0, 1.0), printer=print) -> str:
    acc = list(history.history.get("accuracy", [])) + list(history_fine.history.get("accuracy", []))
    val_acc = list(history.history.get("val_accuracy", [])) + list(history_fine.history.get("val_accuracy", []))
    loss = list(history.history.get("loss", [])) + list(history_fine.history.get("loss", []))
    val_loss = list(history.history.get("val_loss", [])) + list(history_fine.history.get("val_loss", []))
    out_dir = os.path.join(path, "logs")
    os.makedirs(out_dir, exist_ok=True)
    out = os.path.join(out_dir, "plot_dev" + str(num_devices) + ".png")
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        plt.figure(figsize=(8, 8))
        plt.subplot(2, 1, 1)
        plt.plot(acc, label="Training Accuracy")
        plt.plot(val_acc, label="Validation Accuracy")
        if acc_ylim:
            plt.ylim(list(acc_ylim))
        plt.plot([initial_epochs - 1, initial_epochs - 1], plt.ylim(), label="Start Fine Tuning")
        plt.legend(loc="lower right")
        plt.title("Training and Validation Accuracy")
        plt.subplot(2, 1, 2)
        plt.plot(loss, label="Training Loss")
        plt.plot(val_loss, label="Validation Loss")
        if loss_ylim:
            plt.ylim(list(loss_ylim))
        plt.plot([initial_epochs - 1, initial_epochs - 1], plt.ylim(), label="Start Fine Tuning")
        plt.legend(loc="upper right")
        plt.title("Training and Validation Loss")
        plt.xlabel("epoch")
        plt.savefig(out)
        plt.close()
    except ImportError:
        img = np.full((800, 800, 3), 255, dtype=np.uint8)
        n = max(len(acc), 2)
        a_lim = acc_ylim or (min(acc + val_acc + [0.0]), max(acc + val_acc + [1.0]))
        l_lim = loss_ylim or (0.0, max(loss + val_loss + [1.0]))
        _panel(img, 40, 320, [acc, val_acc], a_lim, initial_epochs - 1, n)
        _panel(img, 440, 320, [loss, val_loss], l_lim, initial_epochs - 1, n)
        with open(out, "wb") as f:
            f.write(_png_bytes(img))
    with open(out[:-4] + ".json", "w") as f:
        json.dump({"accuracy": acc, "val_accuracy": val_acc, "loss": loss, "val_loss": val_loss,
                   "start_fine_tuning": initial_epochs - 1, "num_devices": num_devices}, f)
    printer(history.history)
    printer(history_fine.history)
    return out


def write_panels_png(path: str, panels: list, size: int = 800) -> str:
    """Render recorded pyplot-style panels (``{"lines": [(xs, ys, label), ...], "ylim": (lo, hi) | None}``), stacked
    vertically, with the built-in rasteriser. Used by the ``matplotlib.pyplot`` stand-in of
    ``idc_models_b200.run_reference`` when matplotlib is not installed."""
    img = np.full((size, size, 3), 255, dtype=np.uint8)
    n_p = max(len(panels), 1)
    height = (size - 80) // n_p - 40
    colors = [(31, 119, 180), (255, 127, 14), (44, 160, 44), (214, 39, 40)]
    for pi, p in enumerate(panels):
        top = 40 + pi * (height + 80)
        left, right, bottom = 60, size - 20, top + height
        _line(img, left, top, left, bottom, (0, 0, 0))
        _line(img, left, bottom, right, bottom, (0, 0, 0))
        xs_all = [x for xs, _, _ in p["lines"] for x in xs] or [0.0, 1.0]
        ys_all = [y for _, ys, _ in p["lines"] for y in ys] or [0.0, 1.0]
        lo, hi = p.get("ylim") or (min(ys_all), max(ys_all))
        x_lo, x_hi = min(xs_all), max(xs_all)
        xs_span, ys_span = max(x_hi - x_lo, 1e-12), max(hi - lo, 1e-12)
        for li, (xs, ys, _) in enumerate(p["lines"]):
            pts = [(left + (right - left) * ((x - x_lo) / xs_span),
                    bottom - (bottom - top) * ((min(max(y, lo), hi) - lo) / ys_span)) for x, y in zip(xs, ys)]
            for (x0, y0), (x1, y1) in zip(pts[:-1], pts[1:]):
                _line(img, x0, y0, x1, y1, colors[li % len(colors)])
    with open(path, "wb") as f:
        f.write(_png_bytes(img))
    return path

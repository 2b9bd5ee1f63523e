"""PNG outputs of a client run: confusion-matrix heat maps and the local-vs-aggregated bar chart.

Reference ``plot_evaluation`` (client1.py:153-225): ``local_confusion_matrix.png``,
``aggregated_confusion_matrix.png`` (if any) and ``metrics_comparison.png`` in ``clientN_plots/``; 6x6 in heat map
(seaborn ``Blues``, integer annotations, no colour bar), 10x6 in grouped bars (width 0.35, ``#1f77b4`` /
``#ff7f0e``); client 2 saves with ``dpi=300`` (client2.py:155).  Titles follow the CODE (the committed PNGs have the
local/aggregated titles swapped — SURVEY.md Q10).

matplotlib/seaborn are used when importable; this image has neither, so a small Pillow renderer draws the same
figures on the CPU (no graphics API involved anywhere — there is none in the reference either).
"""
from __future__ import annotations

import os
from typing import Optional, Sequence

import numpy as np

LOCAL_COLOR, AGG_COLOR = "#1f77b4", "#ff7f0e"


def _have_matplotlib() -> bool:
    try:
        import matplotlib  # noqa: F401
        import seaborn  # noqa: F401
        return hasattr(matplotlib, "figure")
    except Exception:  # noqa: BLE001
        return False


# ---------------------------------------------------------------------------------------------- Pillow renderer
def _font(size: int):
    from PIL import ImageFont

    try:
        return ImageFont.load_default(size=size)
    except TypeError:  # very old Pillow
        return ImageFont.load_default()


def _blues(t: float):
    # light -> dark blue, close to matplotlib's 'Blues'
    lo, hi = np.array([247, 251, 255.0]), np.array([8, 48, 107.0])
    c = lo + (hi - lo) * float(min(max(t, 0.0), 1.0))
    return tuple(int(v) for v in c)


def _text_center(draw, xy, text, font, fill):
    l, t, r, b = draw.textbbox((0, 0), text, font=font)
    draw.text((xy[0] - (r - l) / 2 - l, xy[1] - (b - t) / 2 - t), text, font=font, fill=fill)


def _pil_confusion(cm, title: str, path: str, dpi: int) -> None:
    from PIL import Image, ImageDraw

    W = H = 6 * dpi
    sc = dpi / 100.0
    img = Image.new("RGB", (W, H), "white")
    d = ImageDraw.Draw(img)
    f_title, f_lab, f_num = _font(int(16 * sc)), _font(int(13 * sc)), _font(int(18 * sc))
    left, top, right, bottom = int(0.13 * W), int(0.12 * H), int(0.95 * W), int(0.88 * H)
    cm = np.asarray(cm, dtype=np.int64).reshape(2, 2)
    vmin, vmax = float(cm.min()), float(cm.max())
    cw, ch = (right - left) / 2, (bottom - top) / 2
    for i in range(2):
        for j in range(2):
            t = (cm[i, j] - vmin) / (vmax - vmin) if vmax > vmin else 0.0
            x0, y0 = left + j * cw, top + i * ch
            d.rectangle([x0, y0, x0 + cw, y0 + ch], fill=_blues(t))
            _text_center(d, (x0 + cw / 2, y0 + ch / 2), str(int(cm[i, j])), f_num, "white" if t > 0.5 else "#262626")
    for j in range(2):
        _text_center(d, (left + (j + 0.5) * cw, bottom + 14 * sc), str(j), f_lab, "black")
        _text_center(d, (left - 14 * sc, top + (j + 0.5) * ch), str(j), f_lab, "black")
    _text_center(d, (W / 2, top / 2), title, f_title, "black")
    _text_center(d, (W / 2, H - 0.04 * H), "Predicted Label", f_lab, "black")
    # vertical y label
    lab = Image.new("RGBA", (int(200 * sc), int(30 * sc)), (255, 255, 255, 0))
    ImageDraw.Draw(lab).text((0, 0), "True Label", font=f_lab, fill="black")
    img.paste(lab.rotate(90, expand=True), (int(0.02 * W), int(H / 2 - 60 * sc)), lab.rotate(90, expand=True))
    img.save(path, dpi=(dpi, dpi))


def _pil_bars(metrics, local_values, agg_values, title: str, path: str, dpi: int) -> None:
    from PIL import Image, ImageDraw

    W, H = 10 * dpi, 6 * dpi
    sc = dpi / 100.0
    img = Image.new("RGB", (W, H), "white")
    d = ImageDraw.Draw(img)
    f_title, f_lab = _font(int(16 * sc)), _font(int(12 * sc))
    left, top, right, bottom = int(0.09 * W), int(0.11 * H), int(0.96 * W), int(0.90 * H)
    vals = list(local_values) + (list(agg_values) if agg_values is not None else [])
    vmax = max(max(vals), 1e-9) * 1.05
    d.rectangle([left, top, right, bottom], outline="black")
    for k in range(6):  # y ticks
        v = vmax * k / 5
        y = bottom - (bottom - top) * k / 5
        d.line([left - 4 * sc, y, left, y], fill="black")
        _text_center(d, (left - 28 * sc, y), f"{v:.0f}" if vmax > 10 else f"{v:.2f}", f_lab, "black")
    n = len(metrics)
    slot = (right - left) / n
    bw = slot * (0.35 if agg_values is not None else 0.8)
    for i, m in enumerate(metrics):
        cx = left + (i + 0.5) * slot
        if agg_values is not None:
            for off, v, col in ((-bw / 2, local_values[i], LOCAL_COLOR), (bw / 2, agg_values[i], AGG_COLOR)):
                y = bottom - (bottom - top) * v / vmax
                d.rectangle([cx + off - bw / 2, y, cx + off + bw / 2, bottom], fill=col)
        else:
            y = bottom - (bottom - top) * local_values[i] / vmax
            d.rectangle([cx - bw / 2, y, cx + bw / 2, bottom], fill=LOCAL_COLOR)
        _text_center(d, (cx, bottom + 14 * sc), m, f_lab, "black")
    _text_center(d, (W / 2, top / 2), title, f_title, "black")
    lx, ly = right - 190 * sc, top + 10 * sc  # legend
    d.rectangle([lx, ly, lx + 22 * sc, ly + 12 * sc], fill=LOCAL_COLOR)
    d.text((lx + 28 * sc, ly - 2 * sc), "Local Model", font=f_lab, fill="black")
    if agg_values is not None:
        d.rectangle([lx, ly + 20 * sc, lx + 22 * sc, ly + 32 * sc], fill=AGG_COLOR)
        d.text((lx + 28 * sc, ly + 18 * sc), "Aggregated Model", font=f_lab, fill="black")
    lab = Image.new("RGBA", (int(120 * sc), int(26 * sc)), (255, 255, 255, 0))
    ImageDraw.Draw(lab).text((0, 0), "Value", font=f_lab, fill="black")
    img.paste(lab.rotate(90, expand=True), (int(0.01 * W), int(H / 2 - 30 * sc)), lab.rotate(90, expand=True))
    img.save(path, dpi=(dpi, dpi))


def _pil_curve(xs, ys, title, xlabel, ylabel, path, dpi, color, diagonal=False, legend="") -> None:
    from PIL import Image, ImageDraw

    W, H = 8 * dpi, 6 * dpi
    sc = dpi / 100.0
    img = Image.new("RGB", (W, H), "white")
    d = ImageDraw.Draw(img)
    f_title, f_lab = _font(int(16 * sc)), _font(int(12 * sc))
    left, top, right, bottom = int(0.1 * W), int(0.1 * H), int(0.95 * W), int(0.88 * H)
    d.rectangle([left, top, right, bottom], outline="black")
    X = lambda v: left + (right - left) * v  # noqa: E731
    Y = lambda v: bottom - (bottom - top) * v / 1.05  # noqa: E731
    if diagonal:
        d.line([X(0), Y(0), X(1), Y(1)], fill="navy", width=max(1, int(2 * sc)))
    pts = [(X(float(a)), Y(float(b))) for a, b in zip(xs, ys)]
    if len(pts) > 1:
        d.line(pts, fill=color, width=max(1, int(2 * sc)))
    _text_center(d, (W / 2, top / 2), title, f_title, "black")
    _text_center(d, (W / 2, H - 0.04 * H), xlabel, f_lab, "black")
    d.text((left + 8 * sc, top + 6 * sc), f"{ylabel}  {legend}", font=f_lab, fill="black")
    img.save(path, dpi=(dpi, dpi))


# ---------------------------------------------------------------------------------------------- public API
def plot_confusion_matrix(cm, title: str, path: str, dpi: int = 100) -> None:
    if _have_matplotlib():
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        import seaborn as sns

        plt.figure(figsize=(6, 6))
        sns.heatmap(np.asarray(cm), annot=True, fmt="d", cmap="Blues", cbar=False)
        plt.title(title)
        plt.ylabel("True Label")
        plt.xlabel("Predicted Label")
        plt.savefig(path, dpi=dpi)
        plt.close()
    else:
        _pil_confusion(cm, title, path, dpi)


def plot_metrics_comparison(local_metrics, aggregated_metrics, title_prefix: str, path: str, dpi: int = 100) -> None:
    names = ["Accuracy", "Precision", "Recall", "F1-Score"]
    pick = lambda m: [m[0], m[2], m[3], m[4]]  # noqa: E731  (accuracy is in percent, the rest fractions: Q8)
    lv = pick(local_metrics)
    av = pick(aggregated_metrics) if aggregated_metrics is not None else None
    title = (f"{title_prefix}: Local vs Aggregated Model Performance" if av is not None
             else f"{title_prefix}: Local Model Performance")
    if _have_matplotlib():
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt

        x = np.arange(len(names))
        fig, ax = plt.subplots(figsize=(10, 6))
        if av is not None:
            ax.bar(x - 0.175, lv, 0.35, label="Local Model", color=LOCAL_COLOR)
            ax.bar(x + 0.175, av, 0.35, label="Aggregated Model", color=AGG_COLOR)
        else:
            ax.bar(x, lv, label="Local Model", color=LOCAL_COLOR)
        ax.set_title(title)
        ax.set_ylabel("Value")
        ax.set_xticks(x)
        ax.set_xticklabels(names)
        ax.legend()
        plt.savefig(path, dpi=dpi)
        plt.close()
    else:
        _pil_bars(names, lv, av, title, path, dpi)


def plot_evaluation(local_metrics, aggregated_metrics=None, output_dir: str = "client1_plots", client_id: int = 1,
                    dpi: Optional[int] = None, extra_curves: bool = False) -> Sequence[str]:
    """Writes the reference's three PNGs (plus optional ROC / PR curves); returns the written paths."""
    os.makedirs(output_dir, exist_ok=True)
    if dpi is None:
        dpi = 100 if client_id == 1 else 300  # client2.py saves with dpi=300, client1.py with the default
    name = f"Client {client_id}"
    written = []
    p = os.path.join(output_dir, "local_confusion_matrix.png")
    plot_confusion_matrix(local_metrics[5], f"{name} Local Model Confusion Matrix", p, dpi)
    written.append(p)
    if aggregated_metrics is not None:
        p = os.path.join(output_dir, "aggregated_confusion_matrix.png")
        plot_confusion_matrix(aggregated_metrics[5], f"{name} Aggregated Model Confusion Matrix", p, dpi)
        written.append(p)
    p = os.path.join(output_dir, "metrics_comparison.png")
    plot_metrics_comparison(local_metrics, aggregated_metrics, name, p, dpi)
    written.append(p)
    if extra_curves:
        from .metrics import pr_curve_points, roc_curve_points

        for tag, m in (("local", local_metrics), ("aggregated", aggregated_metrics)):
            if m is None or not len(m[6]) or len(set(m[6])) < 2:
                continue
            fpr, tpr, auc = roc_curve_points(m[6], m[7])
            p = os.path.join(output_dir, f"{tag}_roc_curve.png")
            _pil_curve(fpr, tpr, f"{name} {tag.title()} Model ROC", "False Positive Rate", "True Positive Rate", p,
                       dpi, "darkorange", diagonal=True, legend=f"(AUC = {auc:.2f})")
            written.append(p)
            pr, rc = pr_curve_points(m[6], m[7])
            p = os.path.join(output_dir, f"{tag}_pr_curve.png")
            _pil_curve(rc, pr, f"{name} {tag.title()} Model Precision-Recall", "Recall", "Precision", p, dpi, "purple")
            written.append(p)
    return written

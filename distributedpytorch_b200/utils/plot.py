"""Image / mask figure (reference: utils/utils.py:38-51 `plot_img_and_mask`, never called there).

There is no display and no graphics stack on the training boxes, so instead of `plt.show()` the figure is
written to a file through matplotlib's headless Agg backend when matplotlib is installed, else as a plain
side-by-side PNG through Pillow."""
from __future__ import annotations

import numpy as np


def plot_img_and_mask(img, mask, path: str = "img_and_mask.png") -> str:
    img = np.asarray(img)
    mask = np.asarray(mask)
    classes = mask.shape[0] if mask.ndim > 2 else 1
    try:
        import matplotlib

        matplotlib.use("Agg")
        import matplotlib.pyplot as plt

        fig, ax = plt.subplots(1, classes + 1)
        ax[0].set_title("Input image")
        ax[0].imshow(img)
        if classes > 1:
            for i in range(classes):
                ax[i + 1].set_title(f"Output mask (class {i + 1})")
                ax[i + 1].imshow(mask[i])
        else:
            ax[1].set_title("Output mask")
            ax[1].imshow(mask)
        for a in ax:
            a.set_xticks([])
            a.set_yticks([])
        fig.savefig(path)
        plt.close(fig)
    except ImportError:
        from PIL import Image

        def to_u8(a):
            a = a.astype(np.float32)
            if a.max() <= 1.0:
                a = a * 255.0
            a = np.clip(a, 0, 255).astype(np.uint8)
            return a if a.ndim == 3 else np.stack([a] * 3, axis=-1)

        panels = [to_u8(img)] + [to_u8(m) for m in (mask if classes > 1 else [mask])]
        h = min(p.shape[0] for p in panels)
        Image.fromarray(np.concatenate([p[:h] for p in panels], axis=1)).save(path)
    return path

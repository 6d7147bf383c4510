"""Image output for developed films: PFM (float32, what `hdrfilm` writes with file_format=pfm,
src/films/hdrfilm.cpp:95-111 / src/libcore/bitmap.cpp write_pfm) and, when OpenCV is importable, EXR."""
import numpy as np


def write_pfm(path, img):
    """img: (H,W,3) or (H,W) float32, top row first (as rendered).  PFM stores rows bottom-up, little endian."""
    img = np.ascontiguousarray(img, np.float32)
    if img.ndim == 3 and img.shape[2] not in (1, 3):
        raise ValueError("PFM needs 1 or 3 channels")
    color = img.ndim == 3 and img.shape[2] == 3
    h, w = img.shape[:2]
    with open(path, "wb") as f:
        f.write(b"PF\n" if color else b"Pf\n")
        f.write(f"{w} {h}\n-1.0\n".encode())
        f.write(np.ascontiguousarray(img[::-1]).astype("<f4").tobytes())


def read_pfm(path):
    with open(path, "rb") as f:
        kind = f.readline().strip()
        w, h = map(int, f.readline().split())
        scale = float(f.readline())
        ch = 3 if kind == b"PF" else 1
        data = np.frombuffer(f.read(), "<f4" if scale < 0 else ">f4", count=w * h * ch)
    img = data.reshape(h, w, ch)[::-1].astype(np.float32)
    return img if ch == 3 else img[:, :, 0]


def write_image(path, img):
    """Dispatch on the extension: .pfm always; .exr needs cv2 (else raises)."""
    if path.lower().endswith(".pfm"):
        return write_pfm(path, img)
    if path.lower().endswith(".exr"):
        import os
        os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
        import cv2
        if not cv2.imwrite(path, np.ascontiguousarray(img[:, :, ::-1], np.float32)):
            raise IOError(f"could not write {path}")
        return None
    raise ValueError(f"unsupported image format: {path}")

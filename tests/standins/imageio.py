"""TEST STAND-IN for imageio (absent from this image).  Not product code.

The reference imports it in src/LBM/utils/io.py:9 and in every demo; the only calls are made AFTER the time loops:
`imageio.get_writer(path, fps=25)`, `writer.append_data(imageio.imread(png))`, `writer.close()`
(demo_2d_LBM.py:127-131).  MP4 encoding is out of scope; the stand-in counts frames and writes nothing."""
import numpy as np

__jrl_standin__ = True


class _Writer:
    def __init__(self, path, **kw):
        self.path, self.frames = path, 0

    def append_data(self, im):
        self.frames += 1

    def close(self):
        pass


def get_writer(path, **kw):
    return _Writer(path, **kw)


def imread(path, **kw):
    try:
        import cv2
        im = cv2.imread(str(path), cv2.IMREAD_UNCHANGED)
        if im is not None:
            return im
    except Exception:
        pass
    return np.zeros((1, 1), np.uint8)

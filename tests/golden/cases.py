"""Seeded end-to-end cases shared by make_golden.py and the tests (name -> kind, w, h, nfeatures, nlevels)."""
from snakeorb import synthetic as S

CASES = {
    "euroc_c1": (("img", 1000), 752, 480, 1000, 8),
    "euroc_c2": (("img", 2000), 752, 480, 1200, 8),
    "kitti_c3": (("img", 3000), 1241, 376, 2000, 8),
    "tum_c4": (("img", 4000), 640, 480, 1000, 8),
    "noise": (("noise", 9), 320, 240, 500, 6),
    "checker": (("checker", 0), 200, 152, 300, 5),
    "constant": (("constant", 0), 200, 152, 300, 5),
    "small4": (("img", 31), 200, 152, 300, 4),
}


def case_image(kind, w, h):
    return S.image(kind[1], w, h) if kind[0] == "img" else S.adversarial(kind[0], w, h, kind[1])

import pathlib
import sys

import numpy as np
import pytest

ROOT = pathlib.Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN = ROOT / 'tests' / 'golden'


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


@pytest.fixture(scope='session')
def golden():
    return {p.stem: np.load(p, allow_pickle=False) for p in GOLDEN.glob('*.npz')}


ANCHOR_SETS = {
    'voc320': ((320, 320), [40, 20, 10], 20),
    'voc640': ((640, 640), [40, 20, 10], 20),
    'tiny128': ((128, 128), [8], 2),
    'stress1120': ((1120, 1120), [140, 70, 35], 80),
    'mid80': ((256, 256), [16, 8], 80),    # C=80  -> 64-row candidate pieces / W=87 encode rows
    'mid100': ((192, 192), [12, 6], 100),  # C=100 -> 32-row candidate pieces / 32-row encode tiles, C % 4 == 0
    'odd7': ((160, 160), [10, 5], 7),      # C=7   -> scalar argmax path, unaligned row widths
}


@pytest.fixture(scope='session')
def oracle_tables():
    from oracle import restated as R
    cache = {}

    def get(name):
        if name not in cache:
            inp, maps, c = ANCHOR_SETS[name]
            cache[name] = R.AnchorTables(inp, maps, R.VOC_PATTERNS, c)
        return cache[name]
    return get


@pytest.fixture(scope='session')
def prior_boxes():
    import object_detector_b200 as odk
    from oracle.restated import VOC_PATTERNS
    cache = {}

    def get(name):
        if name not in cache:
            inp, maps, c = ANCHOR_SETS[name]
            pb = odk.PriorBoxes(inp, maps, c)
            pb.from_dict({'map_sizes': maps, 'pb_size_patterns': VOC_PATTERNS})
            cache[name] = pb
        return cache[name]
    return get


@pytest.fixture
def strict_cap_case():
    """top_k=2 case for layers.py:940: (boxes, classes, confs) — a class-1 box with four class-0 boxes shifted
    by 0.3 of its width around it (IoU 0.538 with it, < 0.45 among themselves) and a far, weak class-0 box."""
    b = np.array([0.3, 0.3, 0.6, 0.6], np.float32)
    d = np.float32(0.09)
    boxes = np.stack([b, b + [d, 0, d, 0], b - [d, 0, d, 0], b + [0, d, 0, d], b - [0, d, 0, d],
                      np.array([0.8, 0.8, 0.95, 0.95], np.float32)]).astype(np.float32)
    classes = np.array([1, 0, 0, 0, 0, 0])
    confs = np.array([0.9, 0.8, 0.7, 0.6, 0.5, 0.4], np.float32)
    return boxes, classes, confs

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

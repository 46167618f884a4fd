"""TEST INFRASTRUCTURE ONLY — loads the UNMODIFIED reference modules, from /root/reference where it is
mounted (the build container), else from the sourceless bytecode ``oracle/make_ref.py`` compiled from those
files into ``oracle/_ref/`` (git-ignored; it travels to the GPU box with the snapshot).

Used by ``oracle/gen_golden.py`` (golden-vector generation), by the ``not gpu`` tests that cross-check
``oracle/restated.py`` against the real reference, and by the ``--impl reference`` arm of ``bench.py``
(the reference's own CPU code as the timed baseline).  Nothing under ``object_detector_b200/`` and
nothing under ``-m gpu`` or ``smoke()`` imports this file.

The reference package cannot be imported whole (``pytoolkit/__init__.py`` pulls
in TensorFlow etc.), so the eight files the detection path needs are loaded by
path under a synthetic ``pytoolkit`` package, with three shims:
``np.bool`` alias, ``sklearn.externals.joblib`` alias and a stub ``tensorflow``
(``PriorBoxes.decode_locs`` imports it unconditionally, od_pb.py:218).
"""
import importlib.machinery
import importlib.util
import pathlib
import sys
import types

import numpy as np

_MOUNTED = pathlib.Path('/root/reference/pytoolkit/pytoolkit')
_COPIED = pathlib.Path(__file__).resolve().parent / '_ref' / 'pytoolkit'
REFERENCE_ROOT = _MOUNTED if (_MOUNTED / 'dl' / 'od_pb.py').is_file() else _COPIED
_FILES = ['utils', 'log', 'math', 'draw', 'ndimage', 'ml', 'dl.backend', 'dl.losses', 'dl.od_pb']  # dl.backend: imported by dl.losses


def _suffix() -> str:
    return '.py' if REFERENCE_ROOT == _MOUNTED else '.rpyc'


def available() -> bool:
    if not (REFERENCE_ROOT / 'dl' / ('od_pb' + _suffix())).is_file():
        return False
    if REFERENCE_ROOT == _COPIED:  # bytecode is tied to the interpreter that compiled it
        import json
        try:
            made_by = json.loads((_COPIED.parent / 'MANIFEST.json').read_text())['python']
        except Exception:  # noqa: BLE001
            return False
        return list(sys.version_info[:2]) == list(made_by)
    return True


def source() -> str:
    """'mounted' (/root/reference), 'compiled' (oracle/_ref) or 'absent'."""
    return 'absent' if not available() else ('mounted' if REFERENCE_ROOT == _MOUNTED else 'compiled')


def load():
    """Returns the synthetic ``pytoolkit`` package (attributes: ml, math, dl.od_pb ...)."""
    if 'pytoolkit' in sys.modules and getattr(sys.modules['pytoolkit'], '_odk_shim', False):
        return sys.modules['pytoolkit']
    if not available():
        raise RuntimeError('reference tree not mounted at %s' % REFERENCE_ROOT)
    sys.dont_write_bytecode = True
    if not hasattr(np, 'bool'):
        np.bool = bool  # ml.py:43,45 use the alias removed in NumPy 1.24
    import joblib
    import sklearn
    ext = types.ModuleType('sklearn.externals')
    ext.joblib = joblib
    sys.modules.setdefault('sklearn.externals', ext)
    sys.modules.setdefault('sklearn.externals.joblib', joblib)
    sklearn.externals = sys.modules['sklearn.externals']
    for missing in ('tensorflow', 'better_exceptions', 'matplotlib', 'matplotlib.pyplot'):
        try:
            __import__(missing)
        except Exception:  # noqa: BLE001 - any import failure gets the stub
            sys.modules[missing] = types.ModuleType(missing)
    pkg = types.ModuleType('pytoolkit')
    pkg.__path__ = [str(REFERENCE_ROOT)]
    pkg._odk_shim = True
    dl = types.ModuleType('pytoolkit.dl')
    dl.__path__ = [str(REFERENCE_ROOT / 'dl')]
    sys.modules['pytoolkit'] = pkg
    sys.modules['pytoolkit.dl'] = dl
    pkg.dl = dl
    for name in _FILES:
        full = 'pytoolkit.' + name
        path = REFERENCE_ROOT / (name.replace('.', '/') + _suffix())
        loader = None if path.suffix == '.py' else importlib.machinery.SourcelessFileLoader(full, str(path))
        spec = importlib.util.spec_from_file_location(full, path, loader=loader)
        mod = importlib.util.module_from_spec(spec)
        sys.modules[full] = mod
        spec.loader.exec_module(mod)
        parent, _, leaf = name.rpartition('.')
        setattr(dl if parent == 'dl' else pkg, leaf, mod)
    return pkg


def load_augment():
    """``load()`` plus the reference's generator / image modules (image.py:233-485), for the bounding-box side of the
    augmentation pipeline.  OpenCV is absent here, so the three PIXEL functions RandomZoom calls (ndimage.resize /
    pad_ltrb / crop) are replaced by shape-only stand-ins; every line of box arithmetic and every RNG draw is the
    reference's own."""
    pkg = load()
    if hasattr(pkg, 'image'):
        return pkg
    for name in ('data_utils', 'generator', 'image'):
        full = 'pytoolkit.' + name
        spec = importlib.util.spec_from_file_location(full, _MOUNTED / (name + '.py'))  # golden generation only: build container
        mod = importlib.util.module_from_spec(spec)
        sys.modules[full] = mod
        spec.loader.exec_module(mod)
        setattr(pkg, name, mod)
    nd = pkg.ndimage
    nd.resize = lambda rgb, width, height, padding=None, interp='lanczos': np.zeros((int(height), int(width), rgb.shape[2]), np.float32)
    nd.pad_ltrb = lambda rgb, x1, y1, x2, y2, padding='edge': np.zeros((rgb.shape[0] + int(y1) + int(y2), rgb.shape[1] + int(x1) + int(x2), rgb.shape[2]), np.float32)
    nd.crop = lambda rgb, x, y, width, height: rgb[int(y):int(y) + int(height), int(x):int(x) + int(width), :]
    return pkg

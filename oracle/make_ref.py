"""TEST / BASELINE INFRASTRUCTURE — compiles the UNMODIFIED reference files of the detection path into
``oracle/_ref/`` so that the reference's own CPU code travels to the GPU box (``/root/reference`` does not
exist there).

    python oracle/make_ref.py          # run in the build container, where /root/reference is mounted

The reference is pure Python, so its "build" is byte-compilation: every file ``oracle/ref_loader.py`` loads
 is compiled FROM WHERE IT LIES under /root/reference with ``py_compile`` into a sourceless
bytecode file (``.rpyc``) under ``oracle/_ref/pytoolkit/`` (same layout), plus a manifest with the SHA-256 of each source.
No reference source text is copied into this repository; ``oracle/_ref/`` is git-ignored but not
gpurun-ignored (like the built ``libodk.so``).  The bytecode is tied to the interpreter version, which is
the same image here and on the GPU box; the manifest records it and ``ref_loader`` refuses a mismatch.
``__graft_entry__.build()`` runs this when ``/root/reference`` is present; ``bench.py --impl reference``
times these modules through the shim loader (``cpu_baseline.kind = "reference"``) and falls back to the
oracle port only if they are absent.
"""
import hashlib
import json
import pathlib
import py_compile
import sys

SRC = pathlib.Path('/root/reference/pytoolkit/pytoolkit')
DST = pathlib.Path(__file__).resolve().parent / '_ref' / 'pytoolkit'
# the eight modules ref_loader names plus dl/backend, which dl/losses imports
SUFFIX = '.rpyc'  # not '.pyc': snapshot tools tend to drop *.pyc; ref_loader loads these files explicitly
MODULES = ['utils', 'log', 'math', 'draw', 'ndimage', 'ml', 'dl/backend', 'dl/losses', 'dl/od_pb']


def make() -> bool:
    """Compiles the modules; returns False (and leaves ``oracle/_ref`` as it is) when the reference tree is not mounted."""
    if not (SRC / 'dl' / 'od_pb.py').is_file():
        return False
    manifest = {}
    for rel in MODULES:
        src, dst = SRC / (rel + '.py'), DST / (rel + SUFFIX)
        dst.parent.mkdir(parents=True, exist_ok=True)
        py_compile.compile(str(src), cfile=str(dst), dfile=str(src), doraise=True,
                           invalidation_mode=py_compile.PycInvalidationMode.UNCHECKED_HASH)
        manifest[rel + '.py'] = hashlib.sha256(src.read_bytes()).hexdigest()
    (DST.parent / 'MANIFEST.json').write_text(json.dumps(
        {'source': str(SRC), 'python': list(sys.version_info[:2]), 'sha256': manifest}, indent=1))
    return True


if __name__ == '__main__':
    print('oracle/_ref written' if make() else f'{SRC} not mounted: nothing compiled')

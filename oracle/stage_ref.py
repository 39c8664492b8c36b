"""Recipe that stages the UNMODIFIED reference model files as oracle/_ref/ (test infrastructure).

The reference (shs2783/Speech-Enhancement) is pure Python with no build step and no package
metadata, so `pip install --target baseline/_ref /root/reference` has nothing to install; this
script is the equivalent: it copies the files the hot path's callers live in -- `models/**.py`,
`losses.py`, `utils.py` -- byte for byte from the reference checkout into the git-ignored
directory `oracle/_ref/` and writes a manifest of their SHA-256 digests.  `oracle/_ref/` is listed
in .gitignore (reference sources never enter this repository's history) but not in .gpurunignore,
so it rides to the GPU box like the git-ignored libse_b200.so.  `__graft_entry__.build()` runs it
when the reference checkout is present; on the GPU box only the staged copy exists.

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs use the staged files (through
oracle/ref_loader.py); nothing under speech_enhancement_b200/ imports them.

    python oracle/stage_ref.py [/root/reference]
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, '_ref')
DEFAULT_SRC = '/root/reference'
TOP_LEVEL = ('losses.py', 'utils.py')


def _files(src):
    out = []
    for base, _, names in os.walk(os.path.join(src, 'models')):
        for n in sorted(names):
            if n.endswith('.py'):
                out.append(os.path.relpath(os.path.join(base, n), src))
    out += [n for n in TOP_LEVEL if os.path.exists(os.path.join(src, n))]
    return sorted(out)


def stage(src=DEFAULT_SRC, dst=DST):
    """copy the reference files; returns the manifest {relative path: sha256}"""
    if not os.path.isdir(os.path.join(src, 'models')):
        raise FileNotFoundError(f'{src}/models not found: no reference checkout to stage')
    if os.path.isdir(dst):
        shutil.rmtree(dst)
    manifest = {}
    for rel in _files(src):
        out = os.path.join(dst, rel)
        os.makedirs(os.path.dirname(out), exist_ok=True)
        shutil.copyfile(os.path.join(src, rel), out)
        with open(out, 'rb') as f:
            manifest[rel] = hashlib.sha256(f.read()).hexdigest()
    with open(os.path.join(dst, 'MANIFEST.json'), 'w') as f:
        json.dump({'source': src, 'files': manifest}, f, indent=1, sort_keys=True)
    return manifest


def verify(dst=DST):
    """True when every staged file still has the digest recorded at staging time (i.e. is unmodified)"""
    try:
        with open(os.path.join(dst, 'MANIFEST.json')) as f:
            files = json.load(f)['files']
    except (OSError, ValueError, KeyError):
        return False
    for rel, digest in files.items():
        try:
            with open(os.path.join(dst, rel), 'rb') as f:
                if hashlib.sha256(f.read()).hexdigest() != digest:
                    return False
        except OSError:
            return False
    return bool(files)


if __name__ == '__main__':
    m = stage(sys.argv[1] if len(sys.argv) > 1 else DEFAULT_SRC)
    print(f'staged {len(m)} reference files into {DST}')

"""Put an unmodified copy of the reference's scripts under baseline/_ref (git-ignored).

`pip install --target baseline/_ref /root/reference` is not possible -- the reference has no setup.py /
pyproject.toml ("Directory '/root/reference' is not installable") -- so "installing" it means copying its flat
scripts byte for byte.  Nothing is edited; `baseline/run_reference.py` runs them as they are.

Two sources, tried in order:
  1. `/root/reference` (the read-only mount on the build machine);
  2. `baseline/reference_scripts.tar.gz` -- an archive of the same six files taken from that mount, kept in the
     repository because a fresh GPU box has neither the mount nor the git-ignored `_ref` directory.
Either way every installed file is checked against the SHA-256 recorded below, so what the reference arm
runs is provably the upstream text.
"""
import hashlib
import os
import shutil
import sys
import tarfile

SRC = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
ARCHIVE = os.path.join(HERE, "reference_scripts.tar.gz")

SHA256 = {
    "cifar10_128batch.py": "58a2305d35f007a2ceddf68a625ac6a9fa1392ab6f9afa2e6a7da82390d7b5c2",
    "cifar10_mpi_mobilenet_224.py": "681b888af69af58b7e26558639ab0cd1a5f5090bb67e6b0c4e5c6cf973d5fadd",
    "cifar10_serial_mobilenet_224.py": "91ec1e178fbd7ca6e12edc87600ab928026bbb19ddeb26f344501fd784f12056",
    "cifar10_gpu_parallel.sh": "230814b2d53a0331cea9b6e3b3ab873d53e1f9df8332196e6986a2d9939c539f",
    "run_gpu128.sh": "964e1b6ae3235bd22870565c43deb75c18b7eba9166220a918b2d5d6e2c47796",
    "serial.slurm": "b584e1f0d3cd4f289e402582c2960d9daef97dd078698029600d6c670e4a4818",
}


def _sha(path: str) -> str:
    with open(path, "rb") as f:
        return hashlib.sha256(f.read()).hexdigest()


def installed(dst: str = DST) -> bool:
    """True when every reference file is present under `dst` with the recorded hash."""
    return all(os.path.isfile(os.path.join(dst, fn)) and _sha(os.path.join(dst, fn)) == h for fn, h in SHA256.items())


def main(dst: str = DST, quiet: bool = False) -> int:
    say = (lambda *a: None) if quiet else print
    if installed(dst):
        say(f"reference already installed in {dst}")
        return 0
    os.makedirs(dst, exist_ok=True)
    if all(os.path.isfile(os.path.join(SRC, fn)) for fn in SHA256):
        for fn in SHA256:
            shutil.copyfile(os.path.join(SRC, fn), os.path.join(dst, fn))
        say(f"copied {len(SHA256)} files from {SRC}")
    elif os.path.isfile(ARCHIVE):
        with tarfile.open(ARCHIVE, "r:gz") as tf:
            for m in tf.getmembers():
                if m.isfile() and os.path.basename(m.name) in SHA256 and os.path.basename(m.name) == m.name:
                    with tf.extractfile(m) as src, open(os.path.join(dst, m.name), "wb") as out:
                        shutil.copyfileobj(src, out)
        say(f"extracted {ARCHIVE}")
    else:
        say(f"neither {SRC} nor {ARCHIVE} present; nothing installed")
        return 1
    bad = [fn for fn, h in SHA256.items() if not os.path.isfile(os.path.join(dst, fn)) or _sha(os.path.join(dst, fn)) != h]
    if bad:
        raise RuntimeError(f"reference install is not byte-identical to upstream: {bad}")
    return 0


if __name__ == "__main__":
    sys.exit(main())

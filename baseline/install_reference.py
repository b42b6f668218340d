"""Put an unmodified copy of the reference's scripts under baseline/_ref (git-ignored).

`pip install --target baseline/_ref /root/reference` is not possible -- the reference has no setup.py /
pyproject.toml ("Directory '/root/reference' is not installable") -- so "installing" it means copying its flat
scripts byte for byte.  Nothing is edited; `baseline/run_reference.py` runs them as they are.
"""
import filecmp
import os
import shutil
import sys

SRC = "/root/reference"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def main() -> int:
    if not os.path.isdir(SRC):
        print(f"{SRC} not present; keeping whatever is already in {DST}")
        return 0
    os.makedirs(DST, exist_ok=True)
    for fn in sorted(os.listdir(SRC)):
        if fn.endswith((".py", ".sh", ".slurm")):
            shutil.copyfile(os.path.join(SRC, fn), os.path.join(DST, fn))
            assert filecmp.cmp(os.path.join(SRC, fn), os.path.join(DST, fn), shallow=False)
            print("copied", fn)
    return 0


if __name__ == "__main__":
    sys.exit(main())

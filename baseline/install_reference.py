"""Install the UNMODIFIED reference under baseline/_ref (git-ignored; travels with gpurun).

1. Try the prescribed ``pip install --no-index --no-build-isolation --find-links /opt/wheelhouse
   --target baseline/_ref /root/reference``.  The reference has neither setup.py nor
   pyproject.toml, so pip refuses ("not installable") -- recorded in DESIGN.md.
2. Fall back to copying the reference's source files byte-for-byte and writing a sha256 manifest
   (baseline/REFERENCE_MANIFEST.json, committed) so bench.py can prove at run time that what it
   imports from baseline/_ref is the unmodified reference.
"""
import hashlib
import json
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference"
DST = os.path.join(HERE, "_ref")
MANIFEST = os.path.join(HERE, "REFERENCE_MANIFEST.json")
FILES = ["main.py", "mpi_tools.py", "models.py", "helpers.py", "utils.py", "data_loader.py",
         "evaluation_pipeline.py", "create_dataset.py", "requirements.txt", "README.md"]


def sha(path):
    with open(path, "rb") as fh:
        return hashlib.sha256(fh.read()).hexdigest()


def main():
    r = subprocess.run([sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links",
                        "/opt/wheelhouse", "--target", DST, REF_SRC], capture_output=True, text=True)
    pip_ok = r.returncode == 0
    print("pip install exit code:", r.returncode, (r.stderr or r.stdout).strip().splitlines()[-1:])
    os.makedirs(DST, exist_ok=True)
    manifest = {"pip_install_ok": pip_ok, "files": {}}
    for f in FILES:
        src = os.path.join(REF_SRC, f)
        if os.path.exists(src):
            shutil.copyfile(src, os.path.join(DST, f))
            manifest["files"][f] = sha(src)
    with open(MANIFEST, "w") as fh:
        json.dump(manifest, fh, indent=1, sort_keys=True)
    print("copied {} files to {}".format(len(manifest["files"]), DST))


if __name__ == "__main__":
    main()

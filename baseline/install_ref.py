"""Place the UNMODIFIED reference under baseline/_ref (git-ignored, travels with gpurun).

`pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target
baseline/_ref /root/reference` fails ("Neither 'setup.py' nor 'pyproject.toml' found": the
reference is three scripts and a CSV, not a package), so "installing" it is a verbatim copy.
"""
import hashlib
import os
import shutil
import sys

SRC = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
FILES = ["FL_CustomMLPCLassifierImplementation_Multiple_Rounds.py", "FL_SkLearn_MLPClassifier_Limitation.py",
         "hyperparameters_tuning.py", "balanced_income_data.csv"]

if __name__ == "__main__":
    os.makedirs(DST, exist_ok=True)
    for f in FILES:
        shutil.copyfile(os.path.join(SRC, f), os.path.join(DST, f))
        print(f, hashlib.sha256(open(os.path.join(DST, f), "rb").read()).hexdigest()[:16])

"""Re-encode a CSV table into the column-wise ``.npz`` that ``fedb200.data`` bundles.

    python tools/pack_income_csv.py /path/to/balanced_income_data.csv \
        federated-learning-with-mpi_b200/data/income_columns.npz
"""
import importlib.util
import os
import sys

_p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                  "federated-learning-with-mpi_b200", "data", "income.py")
_spec = importlib.util.spec_from_file_location("_income", _p)
income = importlib.util.module_from_spec(_spec)
sys.modules["_income"] = income
_spec.loader.exec_module(income)

if __name__ == "__main__":
    src, dst = sys.argv[1], sys.argv[2]
    t = income._read_csv(src)
    income.pack_table(t, dst)
    back = income._read_packed(dst)
    assert back.names == t.names
    for n in t.names:
        assert (back.columns[n] == t.columns[n]).all(), n
    print(f"packed {t.n_rows} rows x {len(t.names)} cols -> {dst} ({os.path.getsize(dst)} bytes)")

"""python tools/full_parity.py [fixture ...] > profiles/r2_full_parity.json
The full-size parity table (K3 / solve_model against the reference-made houv_full*.npz) as JSON."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200 import parity  # noqa: E402

if __name__ == "__main__":
    import contextlib
    names = sys.argv[1:] or ["houv_full", "houv_full_retry"]
    with contextlib.redirect_stdout(sys.stderr):  # solve_model prints its mean errors (houv.py:205)
        reps = [parity.full_parity_report(n) for n in names]
    print(json.dumps(reps, indent=1))

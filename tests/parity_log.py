"""Record of the parity errors the GPU tests ACHIEVE (not only whether they passed): every 100-step / full-size comparison
appends {test, field → rel. L2 error, tolerance, 1-ulp floor of the oracle where measured} to a JSON-lines file.  On the GPU box
the file lands in gpurun_out/ (the only directory that travels back); the committed copy is profiles/r2_parity_errors.jsonl."""
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def record(errors, tol, floor=None, bound=None, note=None):
    path = os.environ.get("NHSTEP_PARITY_LOG", os.path.join(ROOT, "gpurun_out", "parity_errors.jsonl"))
    try:
        os.makedirs(os.path.dirname(path), exist_ok=True)
        rec = {"test": os.environ.get("PYTEST_CURRENT_TEST", "?").split(" ")[0], "tolerance": tol,
               "errors": {k: float(v) for k, v in errors.items()}}
        if floor is not None:
            rec["oracle_1ulp_floor"] = {k: float(v) for k, v in floor.items()}
        if bound is not None:
            rec["bound_applied"] = {k: float(v) for k, v in bound.items()}
        if note:
            rec["note"] = note
        with open(path, "a") as f:
            f.write(json.dumps(rec) + "\n")
    except OSError:
        pass

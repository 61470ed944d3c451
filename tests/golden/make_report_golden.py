"""Golden text for yolo_v1_b200/report.py, produced by the reference's own report code.

The report of src/evaluate.py lives inside its CLI main() (lines 185-290, after the model has been
evaluated), so it cannot be called.  This script cuts exactly those statements out of the UNMODIFIED
file at /root/reference/src/evaluate.py, dedents them and executes them with the names they use bound
to small stand-ins (args, results, dataset, VOCDetectionYOLO.VOC_CLASSES); stdout and the written file
are the golden.  The per-image lines of src/predict.py:94-108 are taken the same way.  The results
dict is the reference's own mAPMetric.compute() on the corr64 case of map_small.npz (stored there).

    PYTHONPATH=/root/reference/src python tests/golden/make_report_golden.py
"""
import contextlib
import io
import json
import os
import sys
import tempfile
import textwrap
import types
from pathlib import Path

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/src"
sys.path.insert(0, REF)
from yolo.dataset import VOCDetectionYOLO  # noqa: E402
from yolo.schemas import BoundingBox, Detection  # noqa: E402


def cut(path, first_marker, last_marker):
    """Lines from the one containing first_marker through the one containing last_marker, dedented."""
    lines = open(path).read().splitlines()
    a = next(i for i, ln in enumerate(lines) if first_marker in ln)
    b = next(i for i, ln in enumerate(lines) if last_marker in ln and i > a)
    return textwrap.dedent("\n".join(lines[a:b + 1]))


def main():
    d = np.load(os.path.join(HERE, "map_small.npz"), allow_pickle=True)
    out = {}
    for case in ("corr64", "ties48"):
        results = {str(k): float(v) for k, v in zip(d[f"{case}/keys"], d[f"{case}/vals"])}
        for k in ("num_small_objects", "num_medium_objects", "num_large_objects"):
            results[k] = int(results[k])          # compute() returns these three as ints
        code = cut(os.path.join(REF, "evaluate.py"), "# Print results", 'print(f"\\nResults saved to {results_file}")')
        with tempfile.TemporaryDirectory() as tmp:
            args = types.SimpleNamespace(num_classes=20, checkpoint=os.path.join(tmp, "best.pth"), datasets=None,
                                         year="2007", image_set="test")
            env = {"results": results, "args": args, "VOCDetectionYOLO": VOCDetectionYOLO, "Path": Path,
                   "dataset": list(range(64 if case == "corr64" else 48)), "years_and_splits": None}
            buf = io.StringIO()
            with contextlib.redirect_stdout(buf):
                exec(compile(code, "evaluate.py[report]", "exec"), env)
            printed = buf.getvalue()
            written = open(os.path.join(tmp, "evaluation_results.txt")).read()
            # the only run-dependent parts: the temporary checkpoint path
            printed = printed.replace(tmp, "<DIR>")
            written = written.replace(tmp, "<DIR>")
        out[case] = {"results": results, "n_images": len(env["dataset"]), "printed": printed, "written": written}

    dets = [Detection(class_id=11, class_name="dog", confidence=0.87654, bbox=BoundingBox(x=0.5, y=0.25, width=0.125, height=0.3333)),
            Detection(class_id=14, class_name="person", confidence=0.5, bbox=BoundingBox(x=0.1, y=0.9, width=1.0, height=0.05))]
    code = cut(os.path.join(REF, "predict.py"), "# Print results", "size ({det.bbox.width:.2f}, {det.bbox.height:.2f})") + "\n        )\n"
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        print("\nProcessing: img/dog.jpg")      # predict.py:94, printed before the model runs
        exec(compile(code, "predict.py[report]", "exec"), {"detections": dets})
    out["predict"] = {"printed": buf.getvalue(),
                      "detections": [[t.class_id, t.class_name, t.confidence, t.bbox.x, t.bbox.y, t.bbox.width, t.bbox.height] for t in dets]}
    with open(os.path.join(HERE, "report.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("wrote report.json")


if __name__ == "__main__":
    main()

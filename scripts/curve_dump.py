"""Dump the GPU 200-step curves next to the reference's (tests/golden/curve_B2_S64_200.json): python scripts/curve_dump.py out.json"""
import json, sys
sys.path[:0] = ["gan-calibrated-semi-supervised-learning_b200", "oracle", "tests"]
import test_curve_gpu as T
gold = json.loads(T.GOLD.read_text())
got = T.gpu_curves(gold["steps"], gold["B"], gold["S"], gold["n_critic"], gold["lambda_gp"])
json.dump(got, open(sys.argv[1], "w"))
for k in T.KEYS:
    a, b = T.windowed(got[k]), T.windowed(gold["curves"][k])
    print(k, "gpu", [round(float(v), 3) for v in a]); print(k, "ref", [round(float(v), 3) for v in b])

"""200-step loss curves of the UNMODIFIED reference modules (build container only; needs /root/reference).
Same driver as oracle/make_golden.py (reference nets + restated loop body with injected refined / dropout masks / GP alpha),
B = 2 patches of 64x64 per step, inputs from cgan_oracle.make_batch(B, S, step).  Only the logged scalars are kept:
tests/golden/curve_B2_S{64,128}_200.json.  TEST INFRASTRUCTURE; run:  PYTHONDONTWRITEBYTECODE=1 python oracle/make_golden_curve.py [steps] [S]"""
import json
import sys
from pathlib import Path

sys.dont_write_bytecode = True
sys.path.insert(0, str(Path(__file__).resolve().parent))
import make_golden as MG  # noqa: E402

if __name__ == "__main__":
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    S = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    rec = MG.run(2, S, steps)
    keys = ("d_loss", "gp", "wd", "loss_G", "loss_iou", "loss_wgan", "g_gradnorm", "d_gradnorm_0", "d_gradnorm_1")
    out = {"B": 2, "S": S, "n_critic": rec["n_critic"], "lambda_gp": rec["lambda_gp"], "steps": steps,
           "curves": {k: [round(s[k], 6) for s in rec["steps"]] for k in keys}}
    dst = Path(__file__).resolve().parent.parent / "tests" / "golden" / f"curve_B2_S{S}_{steps}.json"
    dst.write_text(json.dumps(out))
    print("wrote", dst)

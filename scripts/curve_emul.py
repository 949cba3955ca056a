"""CPU experiment: the oracle's fp32 arithmetic vs the oracle's own bf16-STORAGE emulation over N steps (B=2, S=64).
Shows which logged scalars are sensitive to bf16 storage independent of the CUDA kernels.  python scripts/curve_emul.py [steps]"""
import json, sys
sys.path[:0] = ["oracle", "tests"]
import torch, cgan_oracle as O
torch.set_num_threads(8)
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
KEYS = ("d_loss", "gp", "wd", "loss_G", "loss_iou", "loss_wgan")
st = O.init_state(42)
cur = {k: [] for k in KEYS}
with O.emulate_bf16():
    for i in range(steps):
        b, m, a = O.make_batch(2, 64, i)
        r = O.train_step(st, b, m, a, n_critic=2, lambda_gp=1.0)
        for k in KEYS:
            cur[k].append(float(r[k]))
json.dump(cur, open("gpurun_out/curve_emul.json", "w"))
gold = json.load(open("tests/golden/curve_B2_S64_200.json"))["curves"]
w = lambda v: [round(float(x), 3) for x in torch.tensor(v[:steps], dtype=torch.float64).unfold(0, 20, 20).mean(1)]
for k in KEYS:
    print(k, "emul", w(cur[k])); print(k, "ref ", w(gold[k]))

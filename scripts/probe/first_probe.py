import sys
sys.path[:0] = ["gan-calibrated-semi-supervised-learning_b200"]
import torch
from b200cgan import _lib as L
which, cin, N, S = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
a0 = torch.randn(N, 3, S, S, device="cuda"); a1 = torch.randn(N, 3, S, S, device="cuda") if cin == 6 else None
w = torch.randn(64, cin, 4, 4, device="cuda") * 0.1
y = torch.empty(N, S // 2, S // 2, 64, device="cuda", dtype=torch.bfloat16)
gz = torch.randn(N, S // 2, S // 2, 64, device="cuda").to(torch.bfloat16)
dw = torch.zeros(64, cin, 4, 4, device="cuda"); db = torch.zeros(64, device="cuda")
if which == "fprop":
    L.call("b2c_first_fprop", a0, a1, None, None, None, w, None, None, y, N, S, S, 0.2)
else:
    L.call("b2c_first_wgrad", a0, a1, None, None, None, gz, dw, db, None, None, None, None, N, S, S)
torch.cuda.synchronize()
print(which, cin, N, S, "ok")

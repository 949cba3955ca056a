"""Negative control for tests/test_graph_gpu.py::test_graph_replay_sees_weights_loaded_between_steps: with the trainer's
pre-replay pack refresh suppressed, the step after a mid-run load_state_dict must DIFFER from the eager run."""
import sys
sys.path[:0] = ["gan-calibrated-semi-supervised-learning_b200", "oracle", "tests"]
import torch
import test_graph_gpu as T
from b200cgan import ops, train_step
le, we, _ = T.run(False, 5, reload_at=4)
orig = ops.pack_all
calls = {"n": 0}
def gated(params):
    calls["n"] += 1
    # the trainer's pre-replay refresh is the only pack_all call made OUTSIDE an optimizer step during replays; emulate its
    # absence by skipping every call once the graph exists and we are not capturing
    if torch.cuda.is_current_stream_capturing() or not gated.block:
        return orig(params)
gated.block = False
ops.pack_all = gated
_orig_capture = train_step.CGANTrainer._capture
def cap(self, args):
    _orig_capture(self, args); gated.block = True
train_step.CGANTrainer._capture = cap
lg, wg, _ = T.run(True, 5, reload_at=4)
for k in ("loss_iou", "loss_G", "wd", "gp", "d_loss"):
    print(k, le[4][k], lg[4][k], "WITHIN" if abs(le[4][k] - lg[4][k]) <= 3e-2 * abs(le[4][k]) + 5e-3 else "DIFFERS")

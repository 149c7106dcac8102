"""Host-side scheduling of the CUDA-graph training step (Trainer._graph_step, rad-nerf_b200/trainer.py) with a fake stepper:
which static slot a batch goes to, which kind of graph is replayed around the occupancy updates (every `update_interval` steps
the marched occupancy changes, so nothing may be marched ahead across an update: train_ml.py:173-177), and that the rays staged
for the next step are not copied twice.  Pure host logic: no GPU."""
import radnerf_b200  # noqa: F401
from radnerf_b200.trainer import Trainer


class _FakeStepper:
    def __init__(self):
        self.slots = [("o0", "d0", "t0"), ("o1", "d1", "t1")]
        self.log = []

    def load(self, slot, o=None, d=None, t=None):
        self.log.append(("load", slot, o is not None, d is not None, t is not None))

    def step(self, slot, lr=None, kind="steady"):
        self.log.append(("step", slot, kind))
        return [0.0]


class _FakeEngine:
    def __init__(self):
        self._ahead = {}
        self.eager = 0

    def step(self, o, d, t, lr=None):
        self.eager += 1
        return [0.0]


def _trainer(interval=4):
    tr = object.__new__(Trainer)
    tr.stepper, tr.engine = _FakeStepper(), _FakeEngine()
    tr._slot, tr._staged = 0, None
    tr.global_step, tr.update_interval, tr.lr = 0, interval, 1e-2
    tr._last_results = None
    return tr


def test_kinds_and_slots_around_occupancy_updates():
    tr = _trainer(interval=4)
    nxt = ("o", "d", "t")
    for _ in range(9):
        tr._graph_step("o", "d", "t", nxt)
    steps = [e for e in tr.stepper.log if e[0] == "step"]
    # steps 0..8 with updates before steps 0, 4, 8: first, steady, steady, last | first, steady, steady, last | first
    assert [k for _, _, k in steps] == ["first", "steady", "steady", "last", "first", "steady", "steady", "last", "first"]
    assert [s for _, s, _ in steps] == [0, 1, 0, 1, 0, 1, 0, 1, 0]                 # the two static slots alternate
    assert tr.engine.eager == 0 and tr.global_step == 9


def test_staged_rays_are_not_copied_again():
    tr = _trainer(interval=100)
    tr._graph_step("o", "d", "t", ("o", "d", "t"))            # step 0: loads slot 0 (all three) and stages slot 1 (all three)
    tr._graph_step("o", "d", "t", ("o", "d"))                 # step 1: slot 1 is in place; stages rays only into slot 0
    tr._graph_step("o", "d", "t", ("o", "d", "t"))            # step 2: slot 0 holds the rays, the target is still missing
    loads = [e for e in tr.stepper.log if e[0] == "load"]
    assert loads[0] == ("load", 0, True, True, True) and loads[1] == ("load", 1, True, True, True)
    assert loads[2] == ("load", 0, True, True, False)                          # rays only
    assert loads[3] == ("load", 0, False, False, True)                         # the target follows at its own step
    assert loads[4] == ("load", 1, True, True, True)


def test_without_lookahead_on_a_boundary_the_step_is_eager():
    tr = _trainer(interval=4)
    tr._graph_step("o", "d", "t", None)                        # first step, nothing to march ahead: plain eager step
    assert tr.engine.eager == 1 and [e for e in tr.stepper.log if e[0] == "step"] == []
    tr._graph_step("o", "d", "t", ("o", "d", "t"))            # next step: not staged -> "first"
    assert [e for e in tr.stepper.log if e[0] == "step"] == [("step", 1, "first")]

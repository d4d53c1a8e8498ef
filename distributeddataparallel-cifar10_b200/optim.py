"""`SGD` with the reference's call signature (main.py:27,37,39) and a fused flat update."""
import torch

from .engine import Engine


class SGD(torch.optim.Optimizer):
    """optim.SGD(model.parameters(), lr) -- plain SGD, as the reference configures it (momentum 0,
    weight decay 0).  When the parameters are the 9 tensors of one NetResDeep (views of its flat
    vector) the step is ONE kernel over the flat vector; other parameter sets are refused."""

    def __init__(self, params, lr=1e-3, momentum=0, dampening=0, weight_decay=0, nesterov=False):
        if momentum != 0 or dampening != 0 or weight_decay != 0 or nesterov:
            raise NotImplementedError("the reference's path is plain SGD (main.py:27): momentum/weight decay are not on it")
        super().__init__(params, dict(lr=lr))
        self._engine = None

    def bind(self, engine: Engine):
        self._engine = engine
        return self

    def _find_engine(self):
        if self._engine is not None:
            return self._engine
        for g in self.param_groups:
            for p in g["params"]:
                e = getattr(p, "_nrd_engine", None)
                if e is not None:
                    self._engine = e
                    return e
        raise RuntimeError("optim.SGD is not bound to a NetResDeep engine: construct it with "
                           "`SGD(model.parameters(), lr)` where model is a NetResDeep / DistributedDataParallel from this package "
                           "(their .parameters() carry the binding), or call .bind(model.engine)")

    @torch.no_grad()
    def step(self, closure=None):
        eng = self._find_engine()
        ps = [p for g in self.param_groups for p in g["params"]]
        if eng.flat_params is None or any(p.grad is None for p in ps):
            return None                      # SGD.step skips params without grad
        base = eng.flat_grads.data_ptr()
        if [p.grad.data_ptr() for p in eng.params()] != [base + 4 * o for o in eng.offsets[:9]]:
            raise RuntimeError("gradients are not the engine's flat gradient views; was .grad reassigned?")
        eng.sgd_update(self.param_groups[0]["lr"])
        return None

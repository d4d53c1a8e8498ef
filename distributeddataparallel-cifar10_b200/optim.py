"""`SGD` with the reference's call signature (main.py:27,37,39) and a fused flat update."""
import torch

from .engine import Engine


class SGD(torch.optim.Optimizer):
    """optim.SGD(model.parameters(), lr) -- plain SGD, as the reference's hot path configures it (main.py:27), or with
    momentum as its second script does (ppe_main_ddp.py:133: lr=0.001, momentum=0.9; torch's defaults otherwise: dampening 0,
    no nesterov, no weight decay -- those options are refused).  When the parameters are the 9 tensors of one NetResDeep
    (views of its flat vector) the step is ONE kernel over the flat vector; other parameter sets are refused."""

    def __init__(self, params, lr=1e-3, momentum=0, dampening=0, weight_decay=0, nesterov=False):
        if dampening != 0 or weight_decay != 0 or nesterov or momentum < 0:
            raise NotImplementedError("plain SGD (main.py:27) and SGD with momentum (ppe_main_ddp.py:133) are supported; "
                                      "dampening / nesterov / weight decay are not on the reference's paths")
        super().__init__(params, dict(lr=lr, momentum=momentum))
        self._engine = None
        self._momentum_buf = None

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
        mu = self.param_groups[0].get("momentum", 0)
        if mu == 0:
            eng.sgd_update(self.param_groups[0]["lr"])
            return None
        first = self._momentum_buf is None or self._momentum_buf.device != eng.flat_params.device
        if first:
            self._momentum_buf = torch.zeros_like(eng.flat_params)
        from ._lib import lib, check
        from .engine import _stream_ptr
        check(lib.nrd_sgd_momentum_update(eng.flat_params.data_ptr(), eng.flat_grads.data_ptr(), self._momentum_buf.data_ptr(), eng.flat_params.numel(),
                                          self.param_groups[0]["lr"], mu, 1 if first else 0, _stream_ptr()), "nrd_sgd_momentum_update")
        eng.weights_epoch += 1               # the packed weight copies are re-derived before the next forward (Engine.ensure_packed)
        return None

"""Host logic of tensor_model_parallel_all_reduce (srt/distributed/parallel_state.py): which activations go to the
symmetric-memory all-reduce (b200_tp_all_reduce through B200TPComm.all_reduce) and which stay on torch.distributed —
the decision the reference's GroupCoordinator.all_reduce makes for its custom all-reduce
(python/sglang/srt/distributed/parallel_state.py:463-542).  CPU only: the exchange object is a recording fake."""
import torch

from sglang_npu_b200.srt.distributed import parallel_state as ps


class FakeComm:
    def __init__(self, hidden=8, max_tokens=4, dtype=torch.float32):
        self.hidden, self.max_tokens, self.dtype = hidden, max_tokens, dtype
        self.calls = []
        self._partial = torch.zeros(max_tokens, hidden, dtype=dtype)

    def can_all_reduce(self, x):
        return (x.dtype == self.dtype and x.dim() >= 1 and x.shape[-1] == self.hidden and x.is_contiguous()
                and 0 < x.numel() // self.hidden <= self.max_tokens)

    def partial(self, M):
        return self._partial[:M]

    def all_reduce(self, M):
        self.calls.append(M)
        return self._partial[:M] * 2  # "two ranks holding the same partial"


def _setup(monkeypatch, comm):
    seen = []
    monkeypatch.setattr(ps, "_TP_SIZE", 2)
    monkeypatch.setattr(ps, "_TP_COMM", comm)
    monkeypatch.setattr(ps.dist, "all_reduce", lambda x, op=None, group=None: seen.append(tuple(x.shape)))
    return seen


def test_hidden_sized_activation_takes_the_custom_all_reduce(monkeypatch):
    comm = FakeComm()
    seen = _setup(monkeypatch, comm)
    x = torch.arange(24, dtype=torch.float32).view(3, 1, 8).clone()
    want = x * 2
    out = ps.tensor_model_parallel_all_reduce(x)
    assert out is x and torch.equal(x, want)  # in place, like communication_op.py:11-13
    assert comm.calls == [3] and seen == []


def test_everything_else_stays_on_torch_distributed(monkeypatch):
    comm = FakeComm()
    seen = _setup(monkeypatch, comm)
    for x in (torch.zeros(3, 16), torch.zeros(5, 8), torch.zeros(2, 8, dtype=torch.float64), torch.zeros(8, 2).t(),
              torch.zeros(0, 8)):
        ps.tensor_model_parallel_all_reduce(x)
    assert comm.calls == [] and len(seen) == 5


def test_switch_and_missing_comm_keep_torch_distributed(monkeypatch):
    comm = FakeComm()
    seen = _setup(monkeypatch, comm)
    monkeypatch.setenv("B200_CUSTOM_ALLREDUCE", "0")
    ps.tensor_model_parallel_all_reduce(torch.zeros(2, 8))
    assert comm.calls == [] and len(seen) == 1
    monkeypatch.delenv("B200_CUSTOM_ALLREDUCE")
    monkeypatch.setattr(ps, "_TP_COMM", None)
    ps.tensor_model_parallel_all_reduce(torch.zeros(2, 8))
    assert len(seen) == 2


def test_single_rank_is_the_identity(monkeypatch):
    comm = FakeComm()
    seen = _setup(monkeypatch, comm)
    monkeypatch.setattr(ps, "_TP_SIZE", 1)
    x = torch.ones(2, 8)
    assert ps.tensor_model_parallel_all_reduce(x) is x
    assert comm.calls == [] and seen == []

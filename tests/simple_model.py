"""Toy models shared by the tests (parity: tests/unit/simple_model.py of the reference)."""
import torch


class SimpleModel(torch.nn.Module):

    def __init__(self, hidden_dim=32, nlayers=3, empty_grad=False):
        super().__init__()
        self.linears = torch.nn.ModuleList([torch.nn.Linear(hidden_dim, hidden_dim) for _ in range(nlayers)])
        if empty_grad:
            self.unused = torch.nn.Linear(hidden_dim, hidden_dim)
        self.head = torch.nn.Linear(hidden_dim, 4)
        self.norm = torch.nn.LayerNorm(hidden_dim)

    def forward(self, x, y):
        for l in self.linears:
            x = torch.relu(l(x))
        return torch.nn.functional.cross_entropy(self.head(self.norm(x)).float(), y)


class SimpleFrozenModel(SimpleModel):

    def __init__(self, hidden_dim=32):
        super().__init__(hidden_dim)
        for p in self.linears[0].parameters():
            p.requires_grad = False


def random_batches(n, batch, hidden, seed=0, dtype=torch.float32, device="cpu"):
    g = torch.Generator().manual_seed(seed)
    return [(torch.randn(batch, hidden, generator=g).to(device=device, dtype=dtype), torch.randint(0, 4, (batch, ), generator=g).to(device))
            for _ in range(n)]

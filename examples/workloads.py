"""Synthetic workloads for BASELINE.json's configs (no network: no MNIST download, no
checkpoints).  Shapes follow the reference's examples:
  - MNISTClassifier: examples/ray_ddp_example.py:18-21,167 (layer_1=32, layer_2=64, lr=1e-1,
    batch_size=32) on top of ray.tune.examples.mnist_ptl_mini.LightningMNISTClassifier
    [upstream, restated: 784-l1-l2-10 MLP, log_softmax + NLL, Adam(lr)].
  - BoringModel: ray_lightning/tests/utils.py:24-91 (Linear(32, 2), MSE-to-ones, SGD 0.1).
  - ResNet-50 / BERT-base: BASELINE.json configs[2], configs[3].
"""
import torch
import torch.nn as nn
import torch.nn.functional as F


class MNISTClassifier(nn.Module):
    def __init__(self, config=None):
        super().__init__()
        config = config or {"layer_1": 32, "layer_2": 64, "lr": 1e-1, "batch_size": 32}
        self.config = config
        self.lr = config["lr"]
        self.batch_size = config["batch_size"]
        self.layer_1 = nn.Linear(28 * 28, config["layer_1"])
        self.layer_2 = nn.Linear(config["layer_1"], config["layer_2"])
        self.layer_3 = nn.Linear(config["layer_2"], 10)

    def forward(self, x):
        x = x.view(x.size(0), -1)
        x = torch.relu(self.layer_1(x))
        x = torch.relu(self.layer_2(x))
        return F.log_softmax(self.layer_3(x), dim=1)

    def loss_fn(self, out, y):
        return F.nll_loss(out, y)

    def configure_optimizers(self):
        return torch.optim.Adam(self.parameters(), lr=self.lr)


class BoringNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.layer = nn.Linear(32, 2)

    def forward(self, x):
        return self.layer(x)

    def loss_fn(self, out, y=None):
        return F.mse_loss(out, torch.ones_like(out))

    def configure_optimizers(self):
        return torch.optim.SGD(self.parameters(), lr=0.1)


class _Classifier(nn.Module):
    """Wraps a torchvision-style net with the loss / optimizer the bench uses."""

    def __init__(self, net, lr=0.1, momentum=0.9):
        super().__init__()
        self.net = net
        self.lr, self.momentum = lr, momentum

    def forward(self, x):
        return self.net(x)

    def loss_fn(self, out, y):
        return F.cross_entropy(out, y)

    def configure_optimizers(self):
        return torch.optim.SGD(self.parameters(), lr=self.lr, momentum=self.momentum)


class _BertMLM(nn.Module):
    """BERT masked-LM (BASELINE configs[3]: BERT-base bf16 behind HorovodRayAccelerator): forward
    returns the logits, the loss is token cross-entropy in fp32."""

    def __init__(self, layers=12, lr=1e-4):
        super().__init__()
        from transformers import BertConfig, BertForMaskedLM
        self.net = BertForMaskedLM(BertConfig(num_hidden_layers=layers))
        self.lr = lr

    def forward(self, x):
        return self.net(input_ids=x).logits

    def loss_fn(self, out, y):
        return F.cross_entropy(out.float().view(-1, out.size(-1)), y.view(-1))

    def configure_optimizers(self):
        return torch.optim.AdamW(self.parameters(), lr=self.lr)


def build_model(name, seed=0):
    """Random-init weights of the named architecture, seeded like PL_GLOBAL_SEED
    (ray_ddp.py:154-159 forwards it to every worker)."""
    torch.manual_seed(seed)
    if name == "mnist_mlp":
        return MNISTClassifier()
    if name == "boring":
        return BoringNet()
    if name == "resnet50":
        import torchvision
        return _Classifier(torchvision.models.resnet50())
    if name == "resnet18":
        import torchvision
        return _Classifier(torchvision.models.resnet18())
    if name == "bert_base":
        return _BertMLM(layers=12)
    if name == "bert_tiny":
        return _BertMLM(layers=1)
    raise ValueError(f"unknown workload {name}")


def synthetic_batch(name, batch, rank, step=0, seed=1000):
    """One deterministic CPU batch for (rank, step): generated on the host so CPU and GPU
    arms see identical bits."""
    g = torch.Generator().manual_seed(seed + 7919 * rank + step)
    if name == "mnist_mlp":
        return torch.rand(batch, 1, 28, 28, generator=g), torch.randint(0, 10, (batch,), generator=g)
    if name == "boring":
        return torch.randn(batch, 32, generator=g), None
    if name in ("resnet50", "resnet18"):
        return torch.randn(batch, 3, 224, 224, generator=g), torch.randint(0, 1000, (batch,), generator=g)
    if name in ("bert_base", "bert_tiny"):
        ids = torch.randint(0, 30522, (batch, 128), generator=g)
        return ids, ids.clone()
    raise ValueError(name)


DEFAULT_BATCH = {"mnist_mlp": 32, "boring": 1, "resnet50": 256, "resnet18": 64, "bert_base": 32, "bert_tiny": 4}

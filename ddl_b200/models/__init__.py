"""Small random-init models used only to put a training step behind the loader (stall
measurements, smoke test).  The reference ships no models."""
import torch
from torch import nn


class TinyConvNet(nn.Module):
    def __init__(self, n_classes: int = 10, width: int = 32):
        super().__init__()
        self.net = nn.Sequential(
            nn.Conv2d(3, width, 3, padding=1), nn.ReLU(), nn.MaxPool2d(2),
            nn.Conv2d(width, 2 * width, 3, padding=1), nn.ReLU(), nn.AdaptiveAvgPool2d(1), nn.Flatten(),
            nn.Linear(2 * width, n_classes))

    def forward(self, x):
        return self.net(x)


def resnet50(n_classes: int = 1000) -> nn.Module:
    """torchvision ResNet-50, random init (the BASELINE.json config-2 consumer)."""
    import torchvision

    return torchvision.models.resnet50(weights=None, num_classes=n_classes)


class TinyGPT(nn.Module):
    """A few transformer blocks over token ids (the BASELINE.json config-3 consumer)."""

    def __init__(self, vocab: int = 50257, d: int = 256, layers: int = 2, heads: int = 4, seq: int = 1024):
        super().__init__()
        self.emb = nn.Embedding(vocab, d)
        self.pos = nn.Parameter(torch.zeros(seq, d))
        layer = nn.TransformerEncoderLayer(d, heads, 4 * d, batch_first=True, norm_first=True)
        self.blocks = nn.TransformerEncoder(layer, layers)
        self.head = nn.Linear(d, vocab, bias=False)

    def forward(self, tok):
        x = self.emb(tok) + self.pos[: tok.shape[1]]
        return self.head(self.blocks(x))

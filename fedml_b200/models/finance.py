"""Vertical-FL party models (reference: model/finance/*.py, single_process/classical_vertical_fl/party_models.py:12-121)."""
import torch
import torch.nn as nn


class DenseModel(nn.Module):
    """Linear scorer on a party's local representation; learning rate is owned by the party."""

    def __init__(self, input_dim, output_dim, learning_rate=0.01, bias=True):
        super().__init__()
        self.classifier = nn.Sequential(nn.Linear(input_dim, output_dim, bias=bias))
        self.learning_rate = learning_rate

    def forward(self, x):
        return self.classifier(x)


class LocalModel(nn.Module):
    """Feature extractor of one party."""

    def __init__(self, input_dim, output_dim, learning_rate=0.01):
        super().__init__()
        self.classifier = nn.Sequential(nn.Linear(input_dim, output_dim), nn.LeakyReLU())
        self.output_dim = output_dim
        self.learning_rate = learning_rate

    def forward(self, x):
        return self.classifier(x)

    def get_output_dim(self):
        return self.output_dim

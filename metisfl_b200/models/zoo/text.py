"""Sequence classifier of the reference zoo: embedding(128) -> BiLSTM(64, sequences) -> BiLSTM(64) -> 1 logit
(reference ``examples/keras/models/imdb_lstm.py:20-26``)."""
import torch
import torch.nn as nn
import torch.nn.functional as F

from metisfl_b200.models.model_def import PyTorchDef


class ImdbLSTM(nn.Module, PyTorchDef):
    def __init__(self, max_features: int = 25000, embed: int = 128, hidden: int = 64):
        super().__init__()
        self.embed = nn.Embedding(max_features, embed)
        self.lstm1 = nn.LSTM(embed, hidden, batch_first=True, bidirectional=True)
        self.lstm2 = nn.LSTM(2 * hidden, hidden, batch_first=True, bidirectional=True)
        self.out = nn.Linear(2 * hidden, 1)

    def forward(self, tokens):
        x, _ = self.lstm1(self.embed(tokens.long()))
        _, (h, _) = self.lstm2(x)
        return self.out(torch.cat([h[0], h[1]], dim=-1)).squeeze(-1)

    def loss_fn(self, outputs, targets):
        return F.binary_cross_entropy_with_logits(outputs.float(), targets.float())

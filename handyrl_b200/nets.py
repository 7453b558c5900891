"""Policy/value networks returned by `Environment.net()` — plain PyTorch, outside the CUDA hot path.

Architectures and parameter names follow the reference nets so that checkpoints
(`models/<epoch>.pth`, handyrl/train.py:465-478) load either way:
  SimpleConv2dModel  handyrl/envs/tictactoe.py:18-69   (29,006 parameters)
  GeisterNet         handyrl/envs/geister.py:17-167    (233,832 parameters; ConvLSTM "DRC" body)
Both take batched inputs; GeisterNet additionally accepts the (B, ...) tensors produced
directly on the device by the batched engine.
"""
import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F


# ----------------------------- TicTacToe -----------------------------------
class _ConvBN(nn.Module):
    def __init__(self, cin, cout, k, bn, bias=True):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, k, stride=1, padding=k // 2, bias=bias and not bn)
        self.bn = nn.BatchNorm2d(cout) if bn else None

    def forward(self, x):
        x = self.conv(x)
        return x if self.bn is None else self.bn(x)


class _FlatHead(nn.Module):
    def __init__(self, shape, filters, outputs):
        super().__init__()
        self.flat = shape[1] * shape[2] * filters
        self.conv = _ConvBN(shape[0], filters, 1, bn=False)
        self.activation = nn.LeakyReLU(0.1)
        self.fc = nn.Linear(self.flat, outputs, bias=False)

    def forward(self, x):
        return self.fc(self.activation(self.conv(x)).reshape(-1, self.flat))


class SimpleConv2dModel(nn.Module):
    def __init__(self):
        super().__init__()
        width, depth = 32, 3
        self.conv = nn.Conv2d(3, width, 3, stride=1, padding=1)
        self.blocks = nn.ModuleList([_ConvBN(width, width, 3, bn=True) for _ in range(depth)])
        self.head_p = _FlatHead((width, 3, 3), 2, 9)
        self.head_v = _FlatHead((width, 3, 3), 1, 1)

    def forward(self, x, hidden=None):
        h = F.relu(self.conv(x))
        for blk in self.blocks:
            h = F.relu(blk(h))
        return {'policy': self.head_p(h), 'value': torch.tanh(self.head_v(h))}


# ------------------------------- Geister -------------------------------------
class ConvLSTMCell(nn.Module):
    def __init__(self, input_dim, hidden_dim, kernel_size, bias):
        super().__init__()
        self.input_dim, self.hidden_dim = input_dim, hidden_dim
        self.conv = nn.Conv2d(input_dim + hidden_dim, 4 * hidden_dim, kernel_size,
                              padding=(kernel_size[0] // 2, kernel_size[1] // 2), bias=bias)

    def init_hidden(self, input_size, batch_size):
        if batch_size is None:   # single-sample inference through ModelWrapper: numpy state
            z = np.zeros((self.hidden_dim, *input_size), dtype=np.float32)
            return z, z.copy()
        z = torch.zeros(*batch_size, self.hidden_dim, *input_size)
        return z, z.clone()

    def forward(self, x, state):
        h, c = state
        gates = self.conv(torch.cat([x, h], dim=-3))
        gi, gf, go, gg = torch.split(gates, self.hidden_dim, dim=-3)
        c = torch.sigmoid(gf) * c + torch.sigmoid(gi) * torch.tanh(gg)
        h = torch.sigmoid(go) * torch.tanh(c)
        return h, c


class DRC(nn.Module):
    """Deep repeated ConvLSTM: `num_layers` cells applied `num_repeats` times per step."""

    def __init__(self, num_layers, input_dim, hidden_dim, kernel_size=3, bias=True):
        super().__init__()
        self.blocks = nn.ModuleList([ConvLSTMCell(input_dim, hidden_dim, (kernel_size, kernel_size), bias)
                                     for _ in range(num_layers)])

    def init_hidden(self, input_size, batch_size):
        pairs = [blk.init_hidden(input_size, batch_size) for blk in self.blocks]
        return [p[0] for p in pairs], [p[1] for p in pairs]

    def forward(self, x, hidden, num_repeats):
        if hidden is None:
            hidden = self.init_hidden(x.shape[-2:], x.shape[:-3])
            hidden = ([h.to(x.device) for h in hidden[0]], [c.to(x.device) for c in hidden[1]])
        hs, cs = list(hidden[0]), list(hidden[1])
        for _ in range(num_repeats):
            for i, blk in enumerate(self.blocks):
                hs[i], cs[i] = blk(x, (hs[i], cs[i]))
        return hs[-1], (hs, cs)


class Conv2dHead(nn.Module):
    def __init__(self, shape, filters, out_filters):
        super().__init__()
        self.outputs = shape[1] * shape[2] * out_filters
        self.conv1 = nn.Conv2d(shape[0], filters, 3, stride=1, padding=1, bias=False)
        self.bn = nn.BatchNorm2d(filters)
        self.conv2 = nn.Conv2d(filters, out_filters, 1, bias=False)

    def forward(self, x):
        return self.conv2(F.relu(self.bn(self.conv1(x)))).reshape(-1, self.outputs)


class ScalarHead(nn.Module):
    def __init__(self, shape, filters, outputs):
        super().__init__()
        self.hidden_units = shape[1] * shape[2] * filters
        self.conv = nn.Conv2d(shape[0], filters, 1, bias=False)
        self.bn = nn.BatchNorm2d(filters)
        self.fc = nn.Linear(self.hidden_units, outputs, bias=False)

    def forward(self, x):
        return self.fc(F.relu(self.bn(self.conv(x))).reshape(-1, self.hidden_units))


class GeisterNet(nn.Module):
    def __init__(self):
        super().__init__()
        width, depth, p_filters = 32, 3, 8
        self.input_size = (7 + 18, 6, 6)              # board planes + broadcast scalar features
        self.conv1 = nn.Conv2d(self.input_size[0], width, 3, stride=1, padding=1, bias=False)
        self.bn1 = nn.BatchNorm2d(width)
        self.body = DRC(depth, width, width)
        self.head_p_move = Conv2dHead((2 * width, 6, 6), p_filters, 4)
        self.head_p_set = nn.Linear(1, 70, bias=True)
        self.head_v = ScalarHead((2 * width, 6, 6), 1, 1)
        self.head_r = ScalarHead((2 * width, 6, 6), 1, 1)

    def init_hidden(self, batch_size=None):
        return self.body.init_hidden(self.input_size[1:], batch_size)

    def forward(self, x, hidden):
        board, scalar = x['board'], x['scalar']
        planes = scalar.reshape(*scalar.shape, 1, 1).expand(*scalar.shape, 6, 6)
        h0 = F.relu(self.bn1(self.conv1(torch.cat([planes, board], dim=-3))))
        h, hidden = self.body(h0, hidden, num_repeats=3)
        h = torch.cat([h0, h], dim=-3)
        policy = torch.cat([self.head_p_move(h), self.head_p_set(scalar[:, :1])], dim=-1)
        return {'policy': policy, 'value': torch.tanh(self.head_v(h)), 'return': self.head_r(h), 'hidden': hidden}

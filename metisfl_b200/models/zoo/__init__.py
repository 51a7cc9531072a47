"""Model zoo used by the examples, tests and benchmarks."""
from metisfl_b200.models.zoo.gpt2 import GPT2, gpt2_medium, gpt2_small, lm_loss  # noqa: F401
from metisfl_b200.models.zoo.mlp import FashionMnistMLP, HousingMLP  # noqa: F401
from metisfl_b200.models.zoo.resnet import ResNet18  # noqa: F401
from metisfl_b200.models.zoo.neuro import (AlzheimersDisease2DCNN, AlzheimersDisease3DCNN, BrainAge2DCNN,  # noqa: F401
                                           BrainAge3DCNN)
from metisfl_b200.models.zoo.text import ImdbLSTM  # noqa: F401
from metisfl_b200.models.zoo.vision import CifarCNN, FashionMnistFC, MelanomaFC, MnistFC  # noqa: F401

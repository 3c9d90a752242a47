"""Mirror of ssar/random (random-patch trajectory generator).  DEVICE follows ssar/random/__init__.py:1-3."""
import torch

DEVICE = torch.device("cuda" if torch.cuda.is_available() else "cpu")

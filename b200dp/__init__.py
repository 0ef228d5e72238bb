"""Short alias: ``import b200dp`` is ``import torch_distributed_sandbox_b200``."""
import sys

import torch_distributed_sandbox_b200 as _pkg

sys.modules[__name__] = _pkg

"""b200dp / torch_distributed_sandbox_b200 - a Blackwell-native data-parallel training
framework with the capabilities of AditMeh/torch-distributed-sandbox."""
__version__ = "0.1.0"

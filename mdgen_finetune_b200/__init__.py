"""mdgen_finetune_b200 -- B200-native (sm_100a) implementation of MDGen's denoising hot path:
the LatentMDGenModel velocity network + fixed-grid Euler transport sampler, behind a C ABI."""
from .model import LatentMDGenModel  # noqa: F401
from .transport import Sampler, Transport, create_transport  # noqa: F401
from .prep import prep_batch, decode_samples  # noqa: F401
from .ddp import GradAllReduce, all_reduce_mean_scalar  # noqa: F401

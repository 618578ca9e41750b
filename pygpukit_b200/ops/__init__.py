"""pygpukit_b200.ops -- the operator API of the decode hot path (flat re-export like
pygpukit/ops/basic.py:17-176)."""
from .elementwise import *  # noqa: F401,F403
from .embedding import *  # noqa: F401,F403
from .matmul import *  # noqa: F401,F403
from .moe import *  # noqa: F401,F403
from .nn import *  # noqa: F401,F403
from .sampling import *  # noqa: F401,F403
from .tensor import *  # noqa: F401,F403

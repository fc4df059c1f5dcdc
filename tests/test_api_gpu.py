"""The API tests again, on the real CUDA kernels (same functions, ``gpu`` marker)."""

import pytest

from test_api import *  # noqa: F403
from test_api_models import *  # noqa: F403
from test_parity_api import *  # noqa: F403
from test_fused_k7 import *  # noqa: F403

pytestmark = pytest.mark.gpu

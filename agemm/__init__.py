"""Drop-in module name of the reference's extension (``import agemm``, model/qLinearLayer.py:7-8)."""
from arcquant_b200.agemm import *  # noqa: F401,F403
from arcquant_b200.agemm import (  # noqa: F401
    batch_decode_i4, init_kv_i4, append_kv_i4, batch_decode_f16, init_kv_f16, append_kv_f16)

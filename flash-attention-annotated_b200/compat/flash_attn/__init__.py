"""FA2-style entry points (`flash_attn.flash_attn_interface`) over the fa100 forward.
Put `flash-attention-annotated_b200/compat` in front of the FA2 wheel on sys.path to route an older caller to the B200 kernel."""
from .flash_attn_interface import (  # noqa: F401
    flash_attn_func,
    flash_attn_kvpacked_func,
    flash_attn_qkvpacked_func,
    flash_attn_varlen_func,
    flash_attn_varlen_kvpacked_func,
    flash_attn_varlen_qkvpacked_func,
    flash_attn_with_kvcache,
)

__version__ = "2.8.0+fa100"

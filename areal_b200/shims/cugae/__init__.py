"""Drop-in `cugae` distribution (reference imports it as
`from cugae import cugae1d_nolp_misalign_func`, realhf/impl/model/utils/ppo_functional.py:359-360,
when `pkg_version.is_available("cugae")`, realhf/base/pkg_version.py:7-11 -- hence the dist-info next
to this package).  Put `areal_b200/shims` on PYTHONPATH (or call `areal_b200.patch.install_shims()`)."""
from areal_b200.ppo_functional import gae_1d_nolp_misalign as cugae1d_nolp_misalign_func  # noqa: F401
from areal_b200.ppo_functional import gae_2d_nolp, gae_2d_olp  # noqa: F401

__version__ = "0.1.0"

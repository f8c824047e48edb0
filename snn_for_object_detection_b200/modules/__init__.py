"""Layer generators and custom modules (public surface of the reference's ``models.modules``)."""

from .layer_gen import *  # noqa: F401,F403
from .common import *  # noqa: F401,F403
from .cells import *  # noqa: F401,F403
from .layer_gen import layers_list
from .common import common_list

__all__ = tuple(layers_list) + tuple(common_list)

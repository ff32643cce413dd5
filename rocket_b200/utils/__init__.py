from rocket_b200.utils.collections import apply_to_collection, is_collection  # noqa: F401
from rocket_b200.utils.torch import (  # noqa: F401
    register_default_move_hook, register_move_hook, torch_collate, torch_move)

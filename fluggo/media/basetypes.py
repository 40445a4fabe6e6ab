"""fluggo.media.basetypes for the frame path: v2i, v2f, box2i, rgba (see canvas_b200/basetypes.py)."""
from canvas_b200.basetypes import *          # noqa: F401,F403
from canvas_b200.basetypes import __all__    # noqa: F401

"""fcaf3d_b200 — B200-native hot path of FCAF3D behind the MinkowskiEngine / pcdet_nms operator surfaces."""
import sys

__version__ = "0.1.0"


def install_as_minkowski_engine():
    """Registers fcaf3d_b200.me as the `MinkowskiEngine` module so reference model files import it unchanged."""
    from . import me
    sys.modules["MinkowskiEngine"] = me
    sys.modules["MinkowskiEngine.utils"] = me.utils
    sys.modules["MinkowskiEngine.modules"] = me.modules
    sys.modules["MinkowskiEngine.modules.resnet_block"] = me.modules.resnet_block
    return me

"""baryonification_b200 -- B200-native halo displacement & painting behind the BaryonForge Runners API."""
from . import utils  # noqa: F401
from .utils.io import *  # noqa: F401,F403
from .tables import DisplacementTable, ParamProfileTable, ProfileTable  # noqa: F401
from .pk import PowerSpectrumGrid  # noqa: F401

from . import Runners  # noqa: F401,E402
from .Runners import *  # noqa: F401,F403,E402

__version__ = "0.1.0"

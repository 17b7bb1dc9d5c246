"""Same import surface as /root/reference/part2/part2b/model.py: ``import model as mdl; mdl.VGG11()``."""
import os
import sys

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "../..")))

from ddpb200.models.vgg import VGG11, VGG13, VGG16, VGG19, _VGG, _cfg, _make_layers  # noqa: E402,F401

from ..models.resnet_split import *  # noqa: F401,F403

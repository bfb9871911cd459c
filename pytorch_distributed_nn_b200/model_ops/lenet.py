from ..models.lenet import *  # noqa: F401,F403

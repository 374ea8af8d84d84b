from f110_gym.envs.f110_env import F110Env
from f110_gym.envs.base_classes import Integrator, Simulator
from f110_gym.envs.laser_models import ScanSimulator2D

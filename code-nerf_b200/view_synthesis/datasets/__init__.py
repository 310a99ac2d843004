from .dataset import *

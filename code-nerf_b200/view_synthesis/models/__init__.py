from .model import *

"""thermoextrap_b200: B200-native (sm_100a) implementation of thermoextrap's hot path.

samples -> weighted central comoments -> bootstrap resample -> derivative
expressions -> Taylor / interpolation predict, behind the reference's API.
"""

__version__ = "0.1.0"

"""ramp_b200 -- B200-native state/motion validity checking behind RAMP's Python API."""

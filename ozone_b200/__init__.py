"""ozone_b200 -- B200-native drop-in for the solve-and-differentiate path of LSDOlab/ozone."""
__version__ = '0.1.0'

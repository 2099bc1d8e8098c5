"""B200-native hot path of submarine0418/underwater_image_enhancement."""
__version__ = "0.1.0"

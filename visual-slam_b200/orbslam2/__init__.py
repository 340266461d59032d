"""B200-native drop-in for the front-end modules of p2004dr/visual-slam's `orbslam2`
package (extractor, matcher, initializer, utils): same module and class names, same
signatures and return types; all arithmetic runs in libvslam_b200.so on the GPU."""
__version__ = "0.1.0"

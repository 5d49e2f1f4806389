"""Drop-in name for the reference's Python module: `import pillowfight` gives
the B200 build of ace / gaussian / sobel / canny (libpillowfight_b200).  The
reference's other filters are not part of this path and are not provided."""
from libpillowfight_b200 import (PillowfightB200Error, ace, canny, device_count,  # noqa: F401
                                 gaussian, sobel)

__all__ = ["ace", "gaussian", "sobel", "canny"]

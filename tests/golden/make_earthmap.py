"""Decodes data/earthmap.jpg (copied from the reference's texture/earthmap.jpg) once with PIL and stores a 256x128
RGB8 array, so oracle and GPU read the very same texels and no JPEG decoder is needed at test time."""
import os
import numpy as np
from PIL import Image
root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
im = Image.open(os.path.join(root, "data", "earthmap.jpg")).convert("RGB").resize((256, 128), Image.BILINEAR)
np.save(os.path.join(root, "tests", "golden", "earthmap_256x128.npy"), np.asarray(im, dtype=np.uint8))

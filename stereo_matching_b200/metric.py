"""3-pixel error, as /root/reference/inference.py:185-196 and cal.py:17-28."""
import numpy as np


def three_pixel_error(pred: np.ndarray, gt: np.ndarray) -> float:
    valid_gt_pixels = (gt != 0).astype("float")
    masked_prediction_valid = pred * valid_gt_pixels
    num_valid_gt_pixels = valid_gt_pixels.sum()
    num_error_pixels = (np.abs(masked_prediction_valid - gt) > 3).sum()
    return float(num_error_pixels / num_valid_gt_pixels)

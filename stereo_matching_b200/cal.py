"""Drop-in for /root/reference/cal.py: mean 3-pixel error of results/nyu_post
against results/gt."""
import os
import sys

import cv2

from .metric import three_pixel_error


def run(results="./results", pred_dir="nyu_post"):
    path = os.path.join(results, "gt")
    error = 0.0
    count = 0
    for file in sorted(os.listdir(path)):
        pt = file.split("_")
        disp_targets = cv2.imread(os.path.join(path, file), cv2.IMREAD_GRAYSCALE)
        predicted_map = cv2.imread(os.path.join(results, pred_dir, pt[1] + "_10.png"), cv2.IMREAD_GRAYSCALE)
        if predicted_map is not None and disp_targets.shape == predicted_map.shape:
            error_id = three_pixel_error(predicted_map, disp_targets)
            error += error_id
            print(("%06d" % int(pt[1])) + "_10" + "\t" + ("%f" % error_id))
        count += 1
    print("Mean Error: ", error / count)
    return error / count


if __name__ == "__main__":
    run(*(sys.argv[1:2] or ["./results"]))

"""Prints the stamps of csrc/annotate.cu (h_stamp2: 5x5, h_stamp3: 7x7) from the installed OpenCV: cv2.line with thickness
2 / 3 for a unit step in each chain-code direction, and for a single point.  Bit (dy + r) * (2r + 1) + (dx + r), r = 2 / 3."""
import cv2
import numpy as np

D = [(1, 0), (1, -1), (0, -1), (-1, -1), (-1, 0), (-1, 1), (0, 1), (1, 1)]


def stamp(thickness, code):
    img = np.zeros((11, 11), np.uint8)
    dx, dy = D[code] if code < 8 else (0, 0)
    cv2.line(img, (5, 5), (5 + dx, 5 + dy), 1, thickness)
    bits, r = 0, thickness
    for y, x in zip(*np.nonzero(img)):
        assert abs(int(x) - 5) <= r and abs(int(y) - 5) <= r
        bits |= 1 << ((int(y) - 5 + r) * (2 * r + 1) + (int(x) - 5 + r))
    return bits


if __name__ == "__main__":
    print("cv2", cv2.__version__)
    for t in (2, 3):
        print("thickness", t, ", ".join(hex(stamp(t, c)) + ("ull" if t == 3 else "u") for c in range(9)))

"""Prints the 5x5 stamps of csrc/annotate.cu (c_stamp2 / h_stamp2) from the installed OpenCV: cv2.line with thickness 2
for a unit step in each chain-code direction, and for a single point.  Bit (dy + 2) * 5 + (dx + 2)."""
import cv2
import numpy as np

D = [(1, 0), (1, -1), (0, -1), (-1, -1), (-1, 0), (-1, 1), (0, 1), (1, 1)]


def stamp(thickness, code):
    img = np.zeros((11, 11), np.uint8)
    dx, dy = D[code] if code < 8 else (0, 0)
    cv2.line(img, (5, 5), (5 + dx, 5 + dy), 1, thickness)
    bits = 0
    for y, x in zip(*np.nonzero(img)):
        assert abs(int(x) - 5) <= 2 and abs(int(y) - 5) <= 2
        bits |= 1 << ((int(y) - 5 + 2) * 5 + (int(x) - 5 + 2))
    return bits


if __name__ == "__main__":
    print("cv2", cv2.__version__)
    for c in range(9):
        print(c, hex(stamp(2, c)))

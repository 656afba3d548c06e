"""Constants of envs/sprites.py:5-10 that callers may read."""
WHITE, RED, BLUE, BLACK = (255, 255, 255), (138, 24, 26), (0, 93, 135), (0, 0, 0)
DISP_WIDTH, DISP_HEIGHT = 1200, 800

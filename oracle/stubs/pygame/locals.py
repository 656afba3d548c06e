"""TEST INFRASTRUCTURE ONLY -- the three constants ``from pygame.locals import *`` must supply
(battle_env.py:4, used only inside render())."""
KEYDOWN = 768
K_ESCAPE = 27
QUIT = 256

/* empty stand-in: see SDL.h in this directory */

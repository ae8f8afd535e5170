// placeholder, filled below


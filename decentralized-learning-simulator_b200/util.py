MICROSECONDS = 10 ** 6   # dasklearn/util/__init__.py:1

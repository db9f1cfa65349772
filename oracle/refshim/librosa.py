"""empty stand-in (only imported by the audio transforms)"""

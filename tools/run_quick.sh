python -m pytest tests -m gpu -q -x 2>&1 | tail -3
python tools/profile_render.py bedroom_xml 3 64 2>&1 | grep -E "launch 2"
python tools/profile_render.py bedroom 3 64 2>&1 | grep -E "launch 2"

# rt_render_static: the new test alone in a fresh process, then the whole GPU suite (it is collected last)
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -5
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5

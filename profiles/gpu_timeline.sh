python profiles/timeline.py A 2>&1 | tail -32

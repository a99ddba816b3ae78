"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference quantization bottleneck.

Nothing under ``oracle/`` is product code.  Only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import it, and there
only as the checker (or as the timed CPU baseline), never as the thing shipped.
The product path (``video_tokenizer_b200``) never imports this package.
"""

"""Drop-in replacement of the reference's `style` package (marcinp7/music-style-transfer) whose hot path runs on a
B200 through libmst_b200.so: same module names, function names, positional signatures and return structures for
everything train-model.py and style/style_transfer.py touch.  Host glue only — every grid, every
model forward/backward and every decode goes through the C ABI (include/mst_b200.h); there is no CPU fallback.

Put `music-style-transfer_b200/` on sys.path and `import style` as before.  Documented deviations:
  * grids are float32 CUDA tensors instead of float64 numpy arrays (same shapes; `.shape`, `.sum()`, slicing and
    `prepare_input` behave as train-model.py:99-109 needs);
  * `read_midi` returns an opaque channel collection (the per-message Python objects of the reference never exist);
  * sub-modules of the network cannot be called on their own: the fused kernels implement the four entry points of
    `StyleTransferModel` (forward, extract_style, predict_song_info, apply_style);
  * `ChannelConverter`'s per-note methods and `get_keys_dist` / `get_notes_dist` raise NotImplementedError (the stages
    exist only fused inside the batched kernels); capacity limits: 16 programs-with-notes per MIDI channel, 32 pitched
    channels per decoded song.  The full list is in INTEGRATION.md ("Deviations").
"""

"""tunebfree_b200 -- B200-native batched renderer for tuneBfree's
tonewheel -> scanner vibrato -> Leslie hot path."""

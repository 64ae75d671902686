"""atx — host side of the B200-native Llama backend (engine, scheduler, paged KV, C-ABI binding)."""

from .collectives import (all_gather, axis_index, axis_size, pmean, ppermute, psum, psum_scatter,  # noqa: F401
                          send_recv)

"""``get_chunks`` of kafka/input_output/utils.py:12-40: the reference's only spatial
partitioning primitive (one ``LinearKalman`` per chunk, kafka_test_S2.py:202-205,
kafka_test_Py36.py:241-255).  Kept for drivers written against it; on the GPU a
whole Sentinel-2 tile shard fits one device and ``dist.shard_rows`` partitions
across GPUs instead."""


def get_chunks(nx, ny, block_size=(256, 256)):
    """Yield ``(x_offset, y_offset, nx_valid, ny_valid, chunk_no)`` for the tiles of
    an ``nx`` x ``ny`` image, x-major, ``chunk_no`` counting from 1; edge tiles are
    clipped to the image."""
    bx, by = int(block_size[0]), int(block_size[1])
    if bx <= 0 or by <= 0:
        raise ValueError("block_size must be positive")
    chunk_no = 0
    for x0 in range(0, max(int(nx), 0), bx):
        for y0 in range(0, max(int(ny), 0), by):
            chunk_no += 1
            yield x0, y0, min(bx, nx - x0), min(by, ny - y0), chunk_no

"""iter_parallel (utils/parallel.py:75): one prefetch thread with a queue of depth 1, exceptions forwarded."""
import queue
import threading


def iter_parallel(iterable, depth=1):
    q = queue.Queue(maxsize=depth)
    end = object()

    def work():
        try:
            for item in iterable:
                q.put((item, None))
            q.put((end, None))
        except BaseException as e:      # forwarded to the consumer like ParallelIterable does (:44-51)
            q.put((end, e))

    threading.Thread(target=work, daemon=True).start()
    while True:
        item, err = q.get()
        if err is not None:
            raise err
        if item is end:
            return
        yield item

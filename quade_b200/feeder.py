"""Feeder: one input file streamed into one page-locked buffer by its own thread, ahead of the batches.

The reference pulls one record per file per iteration (src/Quade.py:210-214, :240-243).  Here a batch is a WINDOW of
every file's buffer: [first unconsumed byte, as many bytes as hold about the same number of records in every file).
The thread keeps appending behind the window while the device works, so the main loop never waits for a read it has
only just asked for; what a batch leaves unconsumed (a partial record, or the surplus of a file that is ahead of the
others) simply stays where it is and opens the next window -- no carry-over copy per batch.

The buffer is two halves.  The thread fills one to its end and goes on in the other as soon as the main loop has left
it; the main loop drains a half and, when what is left fits in front of the other half's data (`head` bytes are kept
free there), moves that rest over -- normally a partial record -- and continues in the other half:

    half 0: [ head | data ................ ]      half 1: [ head | data ................ ]
                     rd ->        <- wr (thread)
"""
import threading


class Feeder:
    def __init__(self, reader, array, head, chunk):
        """reader: FastqReader (readinto(arr) -> n, .eof); array: numpy uint8 view of the buffer (two halves); head:
        bytes kept free in front of a half's data, the largest rest that can move over; chunk: bytes per read."""
        self.reader = reader
        self.buf = array
        self.half = int(array.size) // 2
        self.head = int(head)
        assert 0 < self.head < self.half
        self.chunk = max(1, int(chunk))
        self.min_read = max(1, min(self.chunk, self.head // 16))
        self.w_half = 0                     # the half the thread fills, and the end of its data
        self.wr = self.head
        self.r_half = 0                     # the half the windows are in, and the first unconsumed byte
        self.rd = self.head
        self.sealed_end = [0, 0]            # end of the data of a half the thread has left
        self.eof = False                    # the file is exhausted
        self.error = None
        self._stop = False
        self._cv = threading.Condition()
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    # ---------------- reader thread ----------------
    def _run(self):
        try:
            while True:
                with self._cv:
                    while True:
                        if self._stop:
                            return
                        room = self.w_half * self.half + self.half - self.wr
                        if room >= self.min_read:
                            break
                        other = 1 - self.w_half
                        if self.r_half != other:            # the main loop has left the other half: go on there
                            self.sealed_end[self.w_half] = self.wr
                            self.w_half = other
                            self.wr = other * self.half + self.head
                            self._cv.notify_all()
                            continue
                        self._cv.wait()
                    at, n = self.wr, min(self.chunk, room)
                got = self.reader.readinto(self.buf[at:at + n])     # outside the lock: the windows are never touched
                with self._cv:
                    self.wr = at + got
                    if got < n or self.reader.eof:
                        self.eof = True
                    self._cv.notify_all()
                    if self.eof:
                        return
        except BaseException as E:          # noqa: B902 -- surfaced by wait_for() on the consumer side
            with self._cv:
                self.error = E
                self.eof = True
                self._cv.notify_all()

    # ---------------- consumer (main loop) ----------------
    def _end(self):
        """end of the data the windows can reach now, and whether it can still grow"""
        if self.w_half == self.r_half:
            return self.wr, not self.eof
        return self.sealed_end[self.r_half], False

    def wait_for(self, nbytes):
        """Blocks until the window can be nbytes long, or cannot grow (end of the file, or end of the half: the rest
        must be consumed first).  Returns the bytes available."""
        with self._cv:
            while True:
                if self.error is not None:
                    raise self.error
                end, grows = self._end()
                if end - self.rd >= nbytes or not grows:
                    return end - self.rd
                self._cv.wait()

    def window(self, limit, cap=0):
        """(view of the next batch's bytes of this file, True if they end the file).  `limit` bytes at most -- but what is
        left of a half the thread has moved on from goes in ONE window if that is not much more (a third, and no more
        than `cap`): a window that stops just short of the end of the half would leave megabytes to move over, or a
        tiny batch, instead of a partial record."""
        with self._cv:
            end, _ = self._end()
            n = min(end - self.rd, int(limit))
            if self.w_half != self.r_half and end - self.rd <= min(int(cap), int(limit) + int(limit) // 3):
                n = end - self.rd
            last = self.eof and self.w_half == self.r_half and n == end - self.rd
            return self.buf[self.rd:self.rd + n], last

    def consume(self, nbytes):
        """The batch used the first nbytes of the window and its host-to-device copy is complete.  In a half the thread
        has left, a rest that fits in front of the other half's data moves there.  Returns the unconsumed bytes of the
        half."""
        with self._cv:
            self.rd += int(nbytes)
            if self.w_half == self.r_half:
                return self.wr - self.rd
            rest = self.sealed_end[self.r_half] - self.rd
            if rest <= self.head:
                to = self.w_half * self.half + self.head - rest
                if rest:
                    self.buf[to:to + rest] = self.buf[self.rd:self.rd + rest]
                self.rd = to
                self.r_half = self.w_half
                self._cv.notify_all()               # the half left behind is free for the thread
                return self.wr - self.rd
            return rest

    def close(self):
        with self._cv:
            self._stop = True
            self._cv.notify_all()
        self._t.join()

"""Statistics with the reference's trimean definition (bin/statistics.cpp:7-55): (x[n/4] + 2 x[2(n/4)] + x[3(n/4)]) / 4."""
import math


class Statistics:
    def __init__(self):
        self.x = []

    def insert(self, v):
        self.x.append(float(v))

    def clear(self):
        self.x = []

    def count(self):
        return len(self.x)

    def min(self):
        return min(self.x) if self.x else math.nan

    def max(self):
        return max(self.x) if self.x else math.nan

    def avg(self):
        return sum(self.x) / len(self.x) if self.x else math.nan

    def stddev(self):
        if len(self.x) < 2:
            return math.nan
        m = self.avg()
        return math.sqrt(sum((v - m) ** 2 for v in self.x) / (len(self.x) - 1))

    def med(self):
        if not self.x:
            return math.nan
        s = sorted(self.x)
        n = len(s)
        return s[n // 2] if n % 2 else 0.5 * (s[n // 2 - 1] + s[n // 2])

    def trimean(self):
        if not self.x:
            return math.nan
        s = sorted(self.x)
        q = len(s) // 4
        return (s[q] + 2 * s[2 * q] + s[3 * q]) / 4

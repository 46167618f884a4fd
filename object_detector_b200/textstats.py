"""Text output of ``PriorBoxes.check_prior_boxes``: the fixed-width number column and the ``#`` bar histogram that
pytoolkit/pytoolkit/math.py prints (print_histgram, math.py:26-78).  Host-side formatting; the character layout
(column widths, the ``～`` separator, bar scaling, the mean / std trailer) follows the reference so that its log
lines can be compared verbatim."""
import numpy as np

_LINE_WIDTH = 72


def number_column(values, reserve_sign=True):
    """Every value rendered with one shared format: integers right-aligned to the widest magnitude, floats as
    ``%5.3f`` unless all are tiny (< 0.1) or one is >= 10, then ``%8.2e``; one more column when a negative value
    is present (math.py:57-78)."""
    v = np.asarray(values)
    if v.ndim != 1:
        raise AssertionError('a vector is expected')
    extra = int(bool((v < 0).any()) and reserve_sign)
    mag = np.abs(v)
    if np.issubdtype(v.dtype, np.integer) and mag.max() <= 99999999:
        spec = '%dd' % (int(np.ceil(np.log10(mag.max()))) + extra)
    elif bool((mag < 0.1).all()) or bool((mag >= 10).any()):
        spec = '%d.2e' % (8 + extra)
    else:
        spec = '%d.3f' % (5 + extra)
    return [format(x, spec) for x in v]


def histogram_lines(samples, bins=10, value_range=None, weights=None, stats=True, width=_LINE_WIDTH, title=None):
    """The lines of one text histogram (math.py:32-54): optional dashed title, one ``lo～hi (pct%)  ###`` line
    per bin with bars scaled to the fullest bin, then ``mean:`` / ``std:`` of the samples."""
    counts, edges = np.histogram(samples, bins=bins, range=value_range, weights=weights)
    counts = np.asarray(counts)
    labels = number_column(edges)
    share = counts / (counts.sum() + 1e-7)
    fill = counts / (counts.max() + 1e-7)
    pct_width = 5 if bool((share >= 0.995).any()) else 4
    bar_room = width - (2 * (len(labels[0]) + 2) + pct_width + 4)
    out = []
    if title is not None:
        dashes = '-' * ((width - len(title)) // 2 - 1)
        out.append('%s %s %s' % (dashes, title, dashes))
    for k in range(len(counts)):
        pct = format(share[k] * 100, '%d.1f' % pct_width)
        out.append('%s～%s (%s%%)  %s' % (labels[k], labels[k + 1], pct, '#' * int(round(bar_room * fill[k]))))
    if stats:
        out.append('mean: %s' % np.mean(samples))
        out.append('std:  %s' % np.std(samples))
    return out

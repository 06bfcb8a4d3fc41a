"""Text helpers with the reference's function names: character tables of the Shakespeare sets, the TFF tokenisers of the
StackOverflow sets, the LEAF synthetic(alpha, beta) generator.  The packed readers (``data/readers.py``) use the same tables."""

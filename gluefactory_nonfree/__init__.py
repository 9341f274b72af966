"""API-compatible stand-ins for the reference's `gluefactory_nonfree` package (SuperGlue configuration):
`get_model('gluefactory_nonfree.superpoint')` / `get_model('gluefactory_nonfree.superglue')` resolve to the
B200 implementations.  Written from behaviour; no text of the reference's licensed files is used."""

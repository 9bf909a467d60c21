"""TEST INFRASTRUCTURE ONLY -- minimal stand-in for the third-party ``lhotse`` package (absent here,
unpinned in the reference's requirements.txt:6) so that the UNMODIFIED zipvoice/utils/feature.py can be
imported by oracle/make_golden.py.  Only the four names feature.py uses are provided."""

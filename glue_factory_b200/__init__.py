"""B200-native (sm_100a) SuperPoint-open + LightGlue forward path behind the Glue Factory model API."""

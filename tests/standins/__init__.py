"""Duck-typed stand-ins for the two third-party packages at the reference's API boundary that are not in this image
(xarray 0.19, dask): just the attributes `fv3net_b200.labeled.adapt / restore` touch.  Test infrastructure only."""

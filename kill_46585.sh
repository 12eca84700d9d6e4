ssh test@127.0.0.1 "kill -9 1680"
ssh test@127.0.0.1 "kill -9 1681"
ssh test@127.0.0.1 "kill -9 1682"
ssh test@127.0.0.1 "kill -9 1683"
kill -9 1591

ssh test@127.0.0.1 "kill -9 1730"
ssh test@127.0.0.1 "kill -9 1731"
ssh test@127.0.0.1 "kill -9 1732"
ssh test@127.0.0.1 "kill -9 1733"
kill -9 1591

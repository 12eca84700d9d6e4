ssh test@127.0.0.1 "kill -9 2036"
ssh test@127.0.0.1 "kill -9 2037"
ssh test@127.0.0.1 "kill -9 2038"
ssh test@127.0.0.1 "kill -9 2039"
kill -9 1591

ssh test@127.0.0.1 "kill -9 5134"
ssh test@127.0.0.1 "kill -9 5135"
ssh test@127.0.0.1 "kill -9 5136"
ssh test@127.0.0.1 "kill -9 5137"
kill -9 5044
